"""GPU: sustained in-process A/B of one recurrent step (packing + 25 convs + state, device-resident inputs) under several
values of a per-handle tunable: one network handle per value, alternating runs of `steps` steps, several repetitions.

    python tools/ab_step.py --tunable alt_order=0,1 [--shape 4x512x512] [--steps 40] [--reps 4]
"""
import argparse, json, os, sys
from argparse import Namespace
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr
from isosurfacesuperresolution_b200 import synthetic
ap = argparse.ArgumentParser()
ap.add_argument("--tunable", required=True)
ap.add_argument("--shape", default="4x512x512")
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--reps", type=int, default=4)
a = ap.parse_args()
B, h, w = (int(v) for v in a.shape.split("x"))
key, vals = a.tunable.split("=")
vals = [int(v) for v in vals.split(",")]
wts = dict(np.load(os.path.join(ROOT, "tests/golden/enhancenet_weights_unshaded_seed1234.npz")))
low_np, flow_np, _ = synthetic.synthetic_gbuffer(min(B, 2), h, w, 4322, T=2)
low = torch.from_numpy(np.tile(low_np, ((B + 1) // 2, 1, 1, 1, 1))[:B, 1]).cuda().contiguous()
flow = torch.from_numpy(np.tile(flow_np, ((B + 1) // 2, 1, 1, 1, 1))[:B, 0]).cuda().contiguous()
nets = {}
for v in vals:
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(x) for k, x in wts.items()})
    net = net.cuda().eval()
    net._native().set_tunable(key, v)
    nets[v] = net
res = {}
with torch.no_grad():
    prev = isr.initialImage(low, 6, "input", False, 4)
    state = {v: nets[v](low, prev, None, return_state=True)[2] for v in vals}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(a.reps):
        for v in vals:
            net, st = nets[v], state[v]
            for _ in range(3):
                net(low, st, flow, special_mask=True, return_state=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(a.steps):
                net(low, st, flow, special_mask=True, return_state=True)
            e1.record()
            torch.cuda.synchronize()
            res.setdefault("%s=%d" % (key, v), []).append(round(e0.elapsed_time(e1) / a.steps, 4))
print(json.dumps({"shape": [B, h, w], "steps": a.steps, "ms_per_step": res}))
