"""Fixed cost of one trunk-conv launch: back-to-back launches (isr_net_time_conv, CUDA events, PDL on / off) of the first
residual-block conv (fully split) on images from 2 tiles to the 1080p frame's 1020 tiles.  What remains when the tile
work goes to zero is the per-launch floor that the 22 trunk launches of the interactive frame pay."""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); net = net.cuda().eval()
h = net._native()
res = {}
for pdl in (1, 0):
    h.set_tunable("pdl", pdl)
    for shape in ((1, 16, 8), (1, 16, 16), (1, 64, 64), (1, 128, 128), (1, 192, 256), (1, 270, 480), (4, 270, 480)):
        tiles = ((shape[1] + 15) // 16) * ((shape[2] + 7) // 8) * shape[0]
        for prec in ("hybrid", "fp16"):
            ms = min(h.time_conv(*shape, prec, scale=1, iters=200) for _ in range(2))
            res[f"pdl={pdl} {shape} {prec}"] = {"tiles": tiles, "us": ms * 1e3}
            print(f"pdl={pdl} {str(shape):16s} {prec:7s} tiles {tiles:5d}  {ms * 1e3:7.2f} us / launch", flush=True)
h.set_tunable("pdl", 1)
print(json.dumps(res))
