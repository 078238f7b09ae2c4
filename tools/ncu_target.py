#!/usr/bin/env python
"""Minimal target for ncu captures: the recurrent step of one chunk at a given shape, nothing else.

    python tools/ncu_target.py --shape 4x512x512 --frames 3 [--precision hybrid] [--compose]

--compose adds the 12-channel GUI composition of configs[2] (the 1080p interactive frame)."""
import argparse
import os
import sys
from argparse import Namespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr  # noqa: E402
from isosurfacesuperresolution_b200 import ops, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", default="4x512x512")
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--precision", default="hybrid")
ap.add_argument("--compose", action="store_true")
ap.add_argument("--smooth-flow", action="store_true", help="replace the synthetic flow by its noise-free low-frequency part")
ap.add_argument("--tunable", action="append", default=[], help="key=value, process-wide tunable set before the net is built")
args = ap.parse_args()
for kv in args.tunable:
    k, v = kv.split("=")
    isr._lib.set_tunable(k, int(v))
B, h, w = (int(v) for v in args.shape.split("x"))
weights = dict(np.load(os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")))
net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
net = net.cuda().eval()
net.precision = args.precision
shd = isr.ScreenSpaceShading(torch.device("cuda"))
shd.fov(30)
shd.ambient_light_color(np.array([0.1, 0.1, 0.1])); shd.diffuse_light_color(np.array([1.0, 1.0, 1.0]))
shd.specular_light_color(np.array([0.2, 0.2, 0.2])); shd.specular_exponent(16)
shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0]))
shd.ambient_occlusion(0.5); shd.background(np.array([0.0, 0.0, 0.0]))
low_np, flow_np, _ = synthetic.synthetic_gbuffer(1, h, w, 4322, T=2)
if args.smooth_flow:   # the sinusoid of synthetic.synthetic_gbuffer without the per-pixel uniform noise
    yy, xx = np.meshgrid(np.linspace(-1, 1, h, dtype=np.float32), np.linspace(-1, 1, w, dtype=np.float32), indexing="ij")
    for t in range(flow_np.shape[1]):
        flow_np[0, t, 0] = 0.012 * np.sin(2.1 * xx + 0.3 * t) * np.cos(1.3 * yy)
        flow_np[0, t, 1] = 0.012 * np.cos(1.7 * xx) * np.sin(2.3 * yy + 0.2 * t)
low = torch.from_numpy(np.tile(low_np, (B, 1, 1, 1, 1))).cuda()
flow = torch.from_numpy(np.tile(flow_np, (B, 1, 1, 1, 1))).cuda()
render = torch.rand(B, 12, h, w, device="cuda")
with torch.no_grad():
    inp0, inp1, fl0 = low[:, 0].contiguous(), low[:, 1].contiguous(), flow[:, 0].contiguous()
    state = isr.initialImage(inp0, 6, "input", False, 4)
    handle = net._native()
    for _ in range(args.frames):
        _, _, state = handle.forward_recurrent(inp1, state, fl0, True, args.precision, want_raw=False, want_state=True,
                                               out_state=state)
        if args.compose:
            ops.compose(render, state, shd.params22(), True, False, 0.0)
        else:
            shd(state)
torch.cuda.synchronize()
print("ok", args.shape, args.frames)
