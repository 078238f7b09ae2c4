"""GPU: the fused bilinear-upsample + conv launch (postblock.3 -> the 4h x 4w tensor) in its two forms -- four N = 64 phase
convolutions per low-res tile (up2_wide = 0) and two N = 128 phase pairs (up2_wide = 1) -- timed back to back with CUDA
events (isr_net_time_conv scale 3), with the dbg bits that switch the MMAs / the epilogue off.

    python tools/up2_experiments.py [B h w] > gpurun_out/r02_up2_experiments.json
"""
import json, os, sys
from argparse import Namespace
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr
w = dict(np.load(os.path.join(ROOT, "tests/golden/enhancenet_weights_unshaded_seed1234.npz")))
B, hh, ww = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (4, 512, 512)
res = {"shape": [B, hh, ww], "unit": "us per launch (min of 3 x 10 back-to-back launches)"}
for wide in (0, 1):
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    net = net.cuda().eval()
    h = net._native()
    h.set_tunable("up2_wide", wide)
    for dbg, name in ((0, "full"), (32, "full_with_cycle_probe"), (4, "epilogue_handshake_only"), (2, "no_mma")):
        h.set_tunable("dbg", dbg)
        ms = min(h.time_conv(B, hh, ww, "hybrid", scale=3, iters=10) for _ in range(3))
        res["wide%d_%s" % (wide, name)] = round(ms * 1e3, 1)
    h.set_tunable("dbg", 0)
    for dbg, name in ((0, "dominant_conv"), (2048, "dominant_conv_addend_instance"), (2, "dominant_conv_no_mma"), (2050, "dominant_conv_addend_instance_no_mma")):
        h.set_tunable("dbg", dbg)
        res["wide%d_%s" % (wide, name)] = round(min(h.time_conv(B, hh, ww, "hybrid", scale=4, iters=10) for _ in range(3)) * 1e3, 1)
    h.set_tunable("dbg", 0)
print(json.dumps(res, indent=1))
