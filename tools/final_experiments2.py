"""Bounds of the output conv (isr_conv_final2.cuh / isr_conv_final.cuh) on one chunk of configs[4]: full launch, without
the state stores (dbg bit 8), without any global store (bit 6), without MMAs (bit 1), with one issuing warp."""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
res = {}
for fin, groups in ((3, 3), (3, 4), (1, 3), (3, 3), (3, 4)):
    isr._lib.set_tunable("final_stacked", fin)
    isr._lib.set_tunable("final_groups", groups)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); net = net.cuda().eval()
    h = net._native()
    for name, dbg in (("full", 0), ("no_state_stores", 256), ("no_stores", 64), ("no_stores_no_skip", 64 + 128), ("all_math_dummy_stores", 512)):
        h.set_tunable("dbg", dbg)
        res[f"final_stacked={fin} groups={groups} {name}"] = h.time_conv(4, 512, 512, "hybrid", scale=5, iters=20)
        print(f"final_stacked={fin} groups={groups} {name:20s} {res[f'final_stacked={fin} groups={groups} {name}']:.4f} ms", flush=True)
    h.set_tunable("dbg", 0)
isr._lib.set_tunable("final_stacked", 3)
print(json.dumps(res))
