"""A/B of the division-free tile bookkeeping in conv3x3_pair_kernel inside one process (dbg bit 10 = the divisions of the
first form): the first split trunk conv (scale 1) and the dominant 64->64 conv at 4x (scale 4) of one configs[4] chunk."""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); net = net.cuda().eval()
h = net._native()
res = {}
for rep in range(3):
    for dbg in (0, 1024):
        h.set_tunable("dbg", dbg)
        for scale in (1, 4):
            for shape in ((4, 512, 512), (1, 270, 480)):
                ms = h.time_conv(*shape, "hybrid", scale=scale, iters=30)
                res.setdefault(f"dbg={dbg} scale={scale} {shape}", []).append(ms)
h.set_tunable("dbg", 0)
for k, v in res.items():
    print(k, " ".join("%.4f" % x for x in v))
