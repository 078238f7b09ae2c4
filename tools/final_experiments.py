"""GPU: what bounds the output conv (conv3x3_final_kernel)?  dbg bit6 = no global stores, bit7 = no skip loads."""
import json, os, sys
from argparse import Namespace
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import isosurfacesuperresolution_b200 as isr
from isosurfacesuperresolution_b200 import _lib
w = dict(np.load(os.path.join(ROOT, "tests/golden/enhancenet_weights_unshaded_seed1234.npz")))
net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
net = net.cuda().eval()
h = net._native()
res = {}
for fin in (1, 3):
    h.set_tunable("final_issuers", fin)
    for dbg, name in ((0, "full"), (0, "full_again"), (64, "no_stores"), (128, "no_skip_loads"), (192, "no_stores_no_skip")):
        h.set_tunable("dbg", dbg)
        ms = min(h.time_conv(16, 256, 256, "hybrid", scale=5, iters=10) for _ in range(3))
        res[f"issuers{fin}_{name}"] = round(ms * 1e3, 1)
h.set_tunable("dbg", 0); h.set_tunable("final_issuers", 3)
print(json.dumps(res, indent=1))
