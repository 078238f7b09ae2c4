"""GPU: error of every tensor-core operand scheme against the fp32 CPU oracle / golden fixtures, and the time of one
C2-sized step.  Usage (B200): python tools/precision_sweep.py > gpurun_out/precision_sweep.json"""
import json
import os
import sys
from argparse import Namespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr          # noqa: E402
from oracle import oracle as orc                       # noqa: E402  (checker)

G = os.path.join(ROOT, "tests", "golden")
weights = dict(np.load(os.path.join(G, "enhancenet_weights_unshaded_seed1234.npz")))


def net_for(ups="bilinear"):
    n = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample=ups, reconType="residual", useBN=False))
    n.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    return n.cuda().eval()


def err(out, ref):
    out = out.detach().cpu().numpy()
    return [float(np.max(np.abs(out[:, c] - ref[:, c]))) for c in range(ref.shape[1])], orc.psnr(out, ref)


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


res = {}
modes = sys.argv[1:] or ["mixed", "mixed3", "mixed2", "bf16x3", "fp16x3", "fp16", "bf16"]
gf = dict(np.load(os.path.join(G, "enhancenet_forward.npz")))
gs = dict(np.load(os.path.join(G, "sequence_unshaded.npz")))
rng = np.random.RandomState(7)
xc1 = rng.uniform(-1, 1, (1, 101, 128, 128)).astype(np.float32)
ref_c1, _ = orc.enhancenet_forward(weights, xc1, "bilinear", "residual", 5, backend="torch")
# extra seeds for the recurrence (the thin-margin case): 4 frames, 48x64
seqs = []
for seed in (11, 12, 13):
    low, flow, _ = isr.synthetic.synthetic_gbuffer(1, 48, 64, seed, T=4)
    seqs.append((low, flow, orc.unshaded_sequence(weights, low, flow, "bilinear", "input", backend="torch")[0]))
for mk in modes:
    r = {}
    m = mk
    isr._lib.set_tunable("final_w_parts", 0)
    if ":" in mk:                      # "hybrid:K[:F]" = ISR_PREC_HYBRID with K fully split residual blocks, F output-conv weight planes
        parts = mk.split(":")
        m = parts[0]
        isr._lib.set_tunable("trunk_full_blocks", int(parts[1]))
        if len(parts) > 2:
            isr._lib.set_tunable("final_w_parts", int(parts[2]))
    with torch.no_grad():
        for ups in ("bilinear", "nearest"):
            n = net_for(ups)
            n.precision = m
            recon, raw = n(cu(gf["x"]))
            e, p = err(recon, gf[f"recon_{ups}"])
            r[f"golden_{ups}"] = {"max": max(e), "psnr": p}
        n = net_for()
        n.precision = m
        recon, _ = n(cu(xc1))
        e, p = err(recon, ref_c1)
        r["c1_uniform"] = {"max": max(e), "psnr": p}
        runner = isr.SequenceRunner(n, initial_image="input")
        preds, _, _ = runner.run(cu(gs["low"]), cu(gs["flow"]))
        worst = 0.0
        for j in range(3):
            e, p = err(preds[j], gs[f"pred_{j}"])
            worst = max(worst, max(e))
        r["golden_sequence_3f"] = {"max": worst, "psnr_last": p}
        worst = 0.0
        for low, flow, ref in seqs:
            preds, _, _ = runner.run(cu(low), cu(flow))
            for j in range(4):
                e, p = err(preds[j], ref[j])
                worst = max(worst, max(e))
        r["synthetic_sequences_4f_x3"] = {"max": worst}
        # time of one C2-sized forward (16 x 256^2)
        x = torch.randn(16, 101, 256, 256, device="cuda")
        h = n._native()
        for _ in range(2):
            h.forward(x, m, want_raw=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            h.forward(x, m, want_raw=False)
        e1.record()
        torch.cuda.synchronize()
        r["ms_forward_16x256"] = e0.elapsed_time(e1) / 5
    res[mk] = r
    print(mk, json.dumps(r), file=sys.stderr, flush=True)
print(json.dumps(res, indent=1))
