"""GPU: sustained in-process A/B of one conv launch under two values of the `dbg` experiment mask (or of a tunable):
alternating runs of `iters` back-to-back launches (long enough to sit under the power cap), several repetitions.

    python tools/ab_conv.py --scale 4 --dbg 0,2048 [--shape 4x512x512] [--iters 300] [--reps 4] [--tunable up2_wide=0,1]
"""
import argparse, json, os, sys
from argparse import Namespace
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr
ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=int, default=4)
ap.add_argument("--dbg", default="0")
ap.add_argument("--tunable", default="")
ap.add_argument("--shape", default="4x512x512")
ap.add_argument("--iters", type=int, default=300)
ap.add_argument("--reps", type=int, default=4)
a = ap.parse_args()
B, hh, ww = (int(v) for v in a.shape.split("x"))
w = dict(np.load(os.path.join(ROOT, "tests/golden/enhancenet_weights_unshaded_seed1234.npz")))
net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
net = net.cuda().eval()
h = net._native()
variants = [("dbg", int(v)) for v in a.dbg.split(",")]
if a.tunable:
    key, vals = a.tunable.split("=")
    variants = [(key, int(v)) for v in vals.split(",")]
res = {}
for rep in range(a.reps):
    for key, val in variants:
        h.set_tunable(key, val)
        ms = h.time_conv(B, hh, ww, "hybrid", scale=a.scale, iters=a.iters)
        res.setdefault("%s=%d" % (key, val), []).append(round(ms * 1e3, 1))
print(json.dumps({"scale": a.scale, "shape": [B, hh, ww], "iters": a.iters, "us_per_launch": res}))
