#!/usr/bin/env python
"""isr_inpaint_flow (push-pull pyramid on the device) against cv.inpaint(flow, mask == 0, 3, cv.INPAINT_NS), the call of
inference/loadedmodel.py:79-81: different algorithms (parity unpinned), so this records HOW different on the inputs the
path sees: smooth flow fields of the renderer's magnitude (|flow| <= 0.02) with silhouette-shaped holes (everything
outside 3-5 spheres), errors inside the holes relative to the field's range, and both against the true smooth field.

    python tools/inpaint_vs_opencv.py > gpurun_out/r02_inpaint_vs_opencv.json      (GPU box)
"""
import json
import os
import sys

import cv2 as cv
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isosurfacesuperresolution_b200 import ops, synthetic  # noqa: E402

out = {"what": __doc__.split("\n\n")[0], "cases": []}
for (h, w) in ((135, 240), (270, 480)):
    for seed in (1, 2, 3, 4, 5):
        low, _, _ = synthetic.synthetic_gbuffer(1, h, w, seed)
        hole = (low[:, 0, 0] <= 0).astype(np.uint8)
        yy, xx = np.meshgrid(np.linspace(-1, 1, h, dtype=np.float32), np.linspace(-1, 1, w, dtype=np.float32), indexing="ij")
        f = np.stack([0.012 * np.sin(2.1 * xx + seed) * np.cos(1.3 * yy), 0.012 * np.cos(1.7 * xx) * np.sin(2.3 * yy + seed)])[None]
        f = f.astype(np.float32)
        fin = f * (1 - hole[:, None])
        dev = ops.inpaint_flow(torch.from_numpy(fin).cuda(), torch.from_numpy(hole).cuda()).cpu().numpy()
        ref = np.stack([cv.inpaint(fin[0, c], hole[0], 3, cv.INPAINT_NS) for c in (0, 1)])[None]
        inside = np.broadcast_to(hole[:, None] != 0, dev.shape)
        rng = float(f.max() - f.min())
        out["cases"].append({
            "size": [h, w], "seed": seed, "hole_fraction": float(hole.mean()),
            "device_vs_opencv_mean": float(np.abs(dev - ref)[inside].mean() / rng),
            "device_vs_opencv_max": float(np.abs(dev - ref)[inside].max() / rng),
            "device_vs_truth_mean": float(np.abs(dev - f)[inside].mean() / rng),
            "opencv_vs_truth_mean": float(np.abs(ref - f)[inside].mean() / rng),
            "valid_pixels_untouched": bool(np.array_equal(dev[~inside], fin[~inside]))})
for k in ("device_vs_opencv_mean", "device_vs_opencv_max", "device_vs_truth_mean", "opencv_vs_truth_mean"):
    out["worst_" + k] = max(c[k] for c in out["cases"])
out["unit"] = "fraction of the flow field's range, inside the holes"
print(json.dumps(out, indent=1))
