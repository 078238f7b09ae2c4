#!/usr/bin/env python
"""Free-running vs teacher-forced deviation of the 10-frame recurrence (mainVideoUnshaded.py:419-465) from the fp32 CPU
oracle, per frame, for: the reference's own arithmetic on CUDA (stock torch / cuDNN fp32, TF32 off -- the noise floor
SURVEY.md 8(c) asks for) and this library's precision schemes.  Run on the GPU box:

    python tools/recurrence_drift.py > gpurun_out/r02_recurrence_drift.json

Diagnostic tool (not product): imports the oracle as the checker.
"""
import json
import os
import sys
from argparse import Namespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import isosurfacesuperresolution_b200 as isr  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def stats(a, b):
    return {"max_abs": float(np.max(np.abs(a - b))), "psnr_db": orc.psnr(a, b)}


def ours(net, low, flow, prec, teacher=None):
    net.precision = prec
    dev = torch.device("cuda")
    tl, tf = torch.from_numpy(low).to(dev), torch.from_numpy(flow).to(dev)
    preds, state = [], None
    with torch.no_grad():
        for j in range(low.shape[1]):
            inp = tl[:, j].contiguous()
            if j == 0:
                prev = isr.initialImage(inp, 6, "input", False, 4)
                pred, _, state = net(inp, prev, None, return_state=True)
            else:
                if teacher is not None:
                    state = torch.from_numpy(orc.postprocess(teacher[j - 1], unshaded=True)).to(dev)
                pred, _, state = net(inp, state, tf[:, j - 1].contiguous(), special_mask=True, return_state=True)
            preds.append(pred.cpu().numpy())
    return preds


def main():
    weights = dict(np.load(os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")))
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6,
                            Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    net = net.cuda().eval()
    torch.set_num_threads(os.cpu_count() or 1)
    out = {"what": "per-frame deviation from the fp32 CPU oracle over the 10-frame unshaded recurrence, random-init "
                   "reference weights (seed 1234), synthetic G-buffer sequences; free = the run feeds back its own state, "
                   "tf = teacher-forced (frame j computed from the oracle's frame j-1)",
           "torch": torch.__version__, "cudnn": torch.backends.cudnn.version(), "gpu": torch.cuda.get_device_name(0),
           "cases": {}}
    cases = {"golden_16x24": None, "c2_crop_2x256x256": (2, 256, 256, 4322), "c2_crop_2x96x96": (2, 96, 96, 4323)}
    for name, spec in cases.items():
        if spec is None:
            g = dict(np.load(os.path.join(ROOT, "tests", "golden", "sequence_unshaded_10.npz")))
            low, flow = g["low"], g["flow"]
            ref = [g[f"pred_{j}"] for j in range(10)]              # the live reference's own output
        else:
            B, h, w, seed = spec
            low, flow, _ = orc.synthetic_gbuffer(B, h, w, seed, T=10)
            ref = orc.unshaded_sequence_torch(weights, low, flow, "cpu")
        res = {"shape": list(low.shape)}
        runs = {"torch_cuda_fp32": lambda t: orc.unshaded_sequence_torch(weights, low, flow, "cuda", teacher=t)}
        for prec in ("hybrid", "mixed", "bf16x3", "fp16x3"):
            runs[prec] = (lambda p: (lambda t: ours(net, low, flow, p, teacher=t)))(prec)
        for rname, fn in runs.items():
            free, tf = fn(None), fn(ref)
            res[rname] = {"free": [stats(free[j], ref[j]) for j in range(10)],
                          "tf": [stats(tf[j], ref[j]) for j in range(10)]}
        out["cases"][name] = res
        sys.stderr.write(f"{name} done\n")
    print(json.dumps(out, indent=1))
    # compact table on stderr
    for name, res in out["cases"].items():
        for rname in res:
            if rname == "shape":
                continue
            for mode in ("free", "tf"):
                sys.stderr.write(f"{name:20s} {rname:16s} {mode:4s} " +
                                 " ".join("%.1e" % s["max_abs"] for s in res[rname][mode]) + "\n")


if __name__ == "__main__":
    main()
