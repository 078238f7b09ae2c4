"""The reference's interactive loop (mainGUI.py:697-727 render -> :575-608 super-resolution -> display) with every stage on
the device: dense-grid ray-caster (DenseGridRenderer, the DirectRenderer interface) -> LoadedModel.inference on the
renderer's interleaved frame (G-buffer ingest + push-pull flow in-painting + fused recurrent EnhanceNet step) -> shaded
8-bit RGB frame copied to pinned host memory.  An orbiting camera, as CPURenderer/FlowTest does.  Prints one JSON object
with the per-stage and per-frame times (CUDA events, median over the timed frames); measurement only.

    python tools/interactive_loop.py [--frames 60] [--volume 128] [--ao 0|8|32]
"""
import argparse
import json
import math
import os
import sys
from argparse import Namespace

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def blob_volume(n, seed=5):
    rng = np.random.default_rng(seed)
    ax = (np.arange(n) + 0.5) / n - 0.5
    z, y, x = np.meshgrid(ax, ax, ax, indexing="ij")
    v = np.zeros((n, n, n))
    for _ in range(6):
        k = rng.uniform(-9, 9, 3)
        v += rng.uniform(0.3, 1.0) * np.sin(k[0] * x + k[1] * y + k[2] * z + rng.uniform(0, 6.28))
    v -= 6.0 * (x * x + y * y + z * z)
    return (v / np.abs(v).max()).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=60)
    ap.add_argument("--volume", type=int, default=128)
    ap.add_argument("--ao", type=int, default=8)
    ap.add_argument("--low", default="480x270")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        print(json.dumps({"unavailable": "no CUDA device"}))
        return 0
    import isosurfacesuperresolution_b200 as isr
    from isosurfacesuperresolution_b200 import ops
    dev = torch.device("cuda")
    w, h = (int(t) for t in args.low.split("x"))
    r = isr.DenseGridRenderer(dev)
    r.load(blob_volume(args.volume))
    r.send_command("resolution", "%d,%d" % (w, h))
    r.send_command("isovalue", "0.1")
    r.send_command("aosamples", str(args.ao))
    r.send_command("aoradius", "0.05")
    torch.manual_seed(1234)
    net = isr.createNetwork("EnhanceNet", 4, 5 + 6 * 16, [0, 1, 2, 3, 4], 6,
                            Namespace(upsample="bilinear", reconType="residual", useBN=False)).to(dev).eval()
    lm = isr.LoadedModel({"model": net, "parameters": {"initialImage": "input"}}, dev, 4)
    shd = isr.ScreenSpaceShading(dev)
    shd.fov(30)
    shd.ambient_light_color(np.array([0.1, 0.1, 0.1])); shd.diffuse_light_color(np.array([1.0, 1.0, 1.0]))
    shd.specular_light_color(np.array([0.2, 0.2, 0.2])); shd.specular_exponent(16)
    shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0]))
    shd.ambient_occlusion(0.5); shd.background(np.array([0.0, 0.0, 0.0]))
    frame = torch.empty(h, w, 12, device=dev)
    rgb_host = torch.empty((1, 4 * h, 4 * w, 3), dtype=torch.uint8).pin_memory()
    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_render, t_sr, t_show, t_all = [], [], [], []
    prev = None
    warm = 5
    for j in range(args.frames + warm):
        a = 0.02 * j
        r.send_command("cameraOrigin", "%f,%f,%f" % (1.8 * math.sin(a), 0.4, -1.8 * math.cos(a)))
        P = r.params()
        e = [ev() for _ in range(4)]
        e[0].record()
        ops.raycast_iso(r.volume, P, frame, ao_samples=args.ao, hemisphere=r.hemisphere, rotations=r.rotations,
                        binary_steps=r.binary_steps, skip=r.skip)
        r._last = (list(r.origin), list(r.look_at))
        e[1].record()
        prev = lm.inference(frame[None], prev, inpaint="gpu")
        e[2].record()
        rgb8 = ops.shade_rgb8(prev, shd.params22())
        rgb_host.copy_(rgb8, non_blocking=True)
        e[3].record()
        torch.cuda.synchronize()
        if j >= warm:
            t_render.append(e[0].elapsed_time(e[1])); t_sr.append(e[1].elapsed_time(e[2]))
            t_show.append(e[2].elapsed_time(e[3])); t_all.append(e[0].elapsed_time(e[3]))
    med = lambda v: float(np.median(v))
    # the same frame rendered directly at the display resolution (what super-resolution replaces)
    r.send_command("resolution", "%d,%d" % (4 * w, 4 * h))
    big = torch.empty(4 * h, 4 * w, 12, device=dev)
    r.render_direct(big)
    t_full = min(r.render_direct(big) for _ in range(3)) * 1e3
    print(json.dumps({"low_res": [w, h], "display": [4 * w, 4 * h], "volume": args.volume, "ao_samples": args.ao,
                      "frames": args.frames, "ms_render_low": med(t_render), "ms_superres": med(t_sr),
                      "ms_shade_rgb8_d2h": med(t_show), "ms_frame": med(t_all), "fps": 1e3 / med(t_all),
                      "ms_render_at_display_resolution": t_full, "coverage": float(frame[..., 3].mean()),
                      "rgb_mean": float(rgb_host.float().mean())}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
