"""GPU: MMA cadence of the CTA-pair conv with / without concurrent TMA halo loads and epilogue (dbg bits of ConvParams)."""
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
prec = sys.argv[1] if len(sys.argv) > 1 else "hybrid"
for scale in (4,):
    # (bit 3, "MMA warp does not wait for the epilogue", is unsafe with two issuing warps: tmem_full phases can alias)
    for dbg, name in ((32, "full"), (36, "epilogue_handshake_only"), (48, "no_tma_after_fill"),
                      (52, "no_tma_epilogue_handshake_only"), (2, "no_mma")):
        h.set_tunable("dbg", dbg)
        ms = min(h.time_conv(16, 256, 256, prec, scale=scale, iters=10) for _ in range(3))
        res[f"scale{scale}_{name}"] = round(ms * 1e3, 1)
h.set_tunable("dbg", 0)
print(json.dumps(res, indent=1))
