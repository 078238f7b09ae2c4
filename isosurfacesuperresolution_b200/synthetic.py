"""Synthetic workload generator: analytic sphere-scene G-buffers, smooth dense flow, shading parameter
blocks.  Pure numpy and free of any arithmetic of the path itself, so bench.py, the GPU tests and the CPU oracle all
draw bit-identical inputs from here.  (Not part of oracle/: the product bench may import this, never the oracle.)"""
import numpy as np


def synthetic_gbuffer(B, h, w, seed, T=1):
    """Analytic scene of random spheres -> low-res G-buffer sequence.
    Returns low (B,T,5,h,w): mask in [-1,1] (soft edge), unit normals inside, depth in [0.4,0.9];
    flow (B,T,2,h,w) smooth, |flow| <= 0.02; ao (B,T,1,h,w) in [0.3,1]."""
    rng = np.random.RandomState(seed)
    yy, xx = np.meshgrid(np.linspace(-1, 1, h, dtype=np.float32), np.linspace(-1, 1, w, dtype=np.float32), indexing="ij")
    low = np.zeros((B, T, 5, h, w), np.float32)
    flow = np.zeros((B, T, 2, h, w), np.float32)
    ao = np.ones((B, T, 1, h, w), np.float32)
    for b in range(B):
        n_s = rng.randint(3, 6)
        cx, cy = rng.uniform(-0.7, 0.7, n_s), rng.uniform(-0.7, 0.7, n_s)
        rad = rng.uniform(0.15, 0.45, n_s)
        vel = rng.uniform(-0.01, 0.01, (n_s, 2))
        ph = rng.uniform(0, 2 * np.pi, 4)
        for t in range(T):
            mask = np.full((h, w), -1.0, np.float32)
            nrm = np.zeros((3, h, w), np.float32)
            dep = np.zeros((h, w), np.float32)
            zbest = np.full((h, w), -1e9, np.float32)
            for s in range(n_s):
                dx = (xx - (cx[s] + vel[s, 0] * t)) / rad[s]
                dy = (yy - (cy[s] + vel[s, 1] * t)) / rad[s]
                rr = dx * dx + dy * dy
                inside = rr < 1.0
                z = np.sqrt(np.clip(1.0 - rr, 0, None)).astype(np.float32)
                zz = z * rad[s] + 0.1 * s
                upd = inside & (zz > zbest)
                zbest = np.where(upd, zz, zbest)
                edge = np.clip((1.0 - np.sqrt(rr)) * rad[s] * min(h, w) * 0.5, 0, 1).astype(np.float32)
                mask = np.where(upd, edge * 2 - 1, mask)
                nrm = np.where(upd[None], np.stack([dx, -dy, z]).astype(np.float32), nrm)
                dep = np.where(upd, (0.9 - 0.5 * z).astype(np.float32), dep)
            low[b, t, 0], low[b, t, 1:4], low[b, t, 4] = mask, nrm, dep
            flow[b, t, 0] = 0.012 * np.sin(2.1 * xx + ph[0] + 0.3 * t) * np.cos(1.3 * yy + ph[1]) + rng.uniform(-0.004, 0.004, (h, w))
            flow[b, t, 1] = 0.012 * np.cos(1.7 * xx + ph[2]) * np.sin(2.3 * yy + ph[3] + 0.2 * t) + rng.uniform(-0.004, 0.004, (h, w))
            ao[b, t, 0] = np.clip(0.65 + 0.35 * np.sin(5 * xx + ph[0]) * np.cos(4 * yy + ph[2]), 0.3, 1.0)
    return low, flow.astype(np.float32), ao


def shading_params(ambient=(0.1, 0.1, 0.1), diffuse=(1.0, 1.0, 1.0), specular=(0.2, 0.2, 0.2),
                   light=(0.1, 0.1, 1.0), material=(1.0, 0.3, 0.0), background=(0.0, 0.0, 0.0),
                   ao_strength=1.0, exponent=16, enable_specular=True, inverse_ao=False):
    """Packs the 22-float parameter block (same layout as include/isr.h isr_shade)."""
    light = np.asarray(light, np.float64)
    light = light / np.linalg.norm(light)   # shading.py:77 normalises in float64 then casts
    p = np.zeros(22, np.float32)
    p[0:3], p[3:6], p[6:9] = ambient, diffuse, specular
    p[9:12] = light.astype(np.float32)
    p[12:15], p[15:18] = material, background
    p[18], p[19], p[20], p[21] = ao_strength, float(int(exponent)), float(enable_specular), float(inverse_ao)
    return p
