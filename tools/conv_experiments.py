"""GPU: what bounds the dominant conv launch?  Times it with the epilogue stores and / or the MMAs switched off."""
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
for pair in (1, 0):
    h.set_tunable("cta_pair", pair)
    for scale in (4, 1):
        for dbg, name in ((0, "full"), (1, "no_stores"), (2, "no_mma"), (6, "no_mma_no_epilogue"), (4, "no_epilogue"), (14, "loads_only_decoupled"), (8, "mma_not_waiting_epilogue")):
            if dbg & 10 and not pair:
                continue
            h.set_tunable("dbg", dbg)
            ms = min(h.time_conv(16, 256, 256, "mixed", scale=scale, iters=10) for _ in range(3))
            res[f"pair{pair}_scale{scale}_{name}"] = round(ms * 1e3, 1)
h.set_tunable("dbg", 0); h.set_tunable("cta_pair", 1)
print(json.dumps(res, indent=1))
