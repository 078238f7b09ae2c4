"""Do two independent chunks of configs[4] on two streams overlap usefully (memory-bound packing / shading of one chunk under
the tensor-bound convs of the other)?  Device-resident recurrent steps of 8 chunks, one stream vs two streams (each stream
its own network handle / workspace), CUDA events."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
def mk():
    net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); return net.cuda().eval()
nets = [mk(), mk()]
hs = [n._native() for n in nets]
shd = isr.ScreenSpaceShading(torch.device("cuda"))
shd.fov(30); shd.ambient_light_color(np.array([0.1]*3)); shd.diffuse_light_color(np.array([1.0]*3)); shd.specular_light_color(np.array([0.2]*3)); shd.specular_exponent(16)
shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0])); shd.ambient_occlusion(0.5); shd.background(np.array([0.0]*3))
B, hh, ww, NCH = 4, 512, 512, 8
low = torch.rand(B, 5, hh, ww, device="cuda"); flow = (torch.rand(B, 2, hh, ww, device="cuda") - 0.5) * 0.04
states = [torch.rand(B, 6, 4 * hh, 4 * ww, device="cuda") for _ in range(NCH)]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
def run(nstreams, steps):
    for _ in range(steps):
        for c in range(NCH):
            k = c % nstreams
            with torch.cuda.stream(streams[k]):
                _, _, st = hs[k].forward_recurrent(low, states[c], flow, True, "hybrid", want_raw=False, want_state=True, out_state=states[c])
                shd(st)
with torch.no_grad():
    for rep in range(3):
        for ns in (1, 2):
            run(ns, 1); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for s in streams[:ns]: s.wait_event(e0)
            run(ns, 4)
            for s in streams[:ns]: torch.cuda.current_stream().wait_stream(s)
            e1.record(); torch.cuda.synchronize()
            print(f"streams={ns}: {e0.elapsed_time(e1) / (4 * NCH):.3f} ms per chunk step", flush=True)
