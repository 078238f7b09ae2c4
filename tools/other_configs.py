"""Device-resident throughput of the other BASELINE.json configs (parity-test cases, not bench lines): one recurrent step
of configs[3] (shaded 56 -> 3, 4 x 540x960 -> 2160x3840) and of one 4-sequence chunk of configs[4] (unshaded,
512^2 -> 2048^2).  CUDA events, 3 warm-up + 10 timed steps.  Prints one JSON object."""
import json
import os
import sys
from argparse import Namespace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import isosurfacesuperresolution_b200 as isr   # noqa: E402

opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
res = {}
for name, B, h, w, c_low, c_prev, mask in (("configs[3] shaded 4 x 540x960", 4, 540, 960, 8, 3, [0, 1, 2]),
                                           ("configs[4] chunk 4 x 512x512", 4, 512, 512, 5, 6, [0, 1, 2, 3, 4])):
    torch.manual_seed(1234)
    net = isr.createNetwork("EnhanceNet", 4, c_low + 16 * c_prev, mask, c_prev, opt).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(5)
    low = torch.rand(B, c_low, h, w, device="cuda", generator=g)
    flow = (torch.rand(B, 2, h, w, device="cuda", generator=g) - 0.5) * 0.04
    state = torch.rand(B, c_prev, 4 * h, 4 * w, device="cuda", generator=g)
    with torch.no_grad():
        for _ in range(3):
            pred, _, state = net(low, state, flow, special_mask=(c_prev == 6), return_state=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            pred, _, state = net(low, state, flow, special_mask=(c_prev == 6), return_state=True)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    res[name] = {"ms_per_step": ms, "mpix_per_s": B * 16 * h * w / ms / 1e3, "finite": bool(torch.isfinite(pred).all())}
    del net, low, flow, state, pred
    torch.cuda.empty_cache()
print(json.dumps(res))
