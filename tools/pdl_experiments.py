"""A/B of programmatic dependent launch for the small kernels (tunable pdl_small) inside one process: the device-resident
recurrent step of one configs[4] chunk (CUDA events over 30 steps) and the 1080p frame as a CUDA graph (p50 of 200)."""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
from isosurfacesuperresolution_b200 import ops, synthetic
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); net = net.cuda().eval()
h = net._native()
shd = isr.ScreenSpaceShading(torch.device("cuda"))
shd.fov(30); shd.ambient_light_color(np.array([0.1]*3)); shd.diffuse_light_color(np.array([1.0]*3)); shd.specular_light_color(np.array([0.2]*3)); shd.specular_exponent(16)
shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0])); shd.ambient_occlusion(0.5); shd.background(np.array([0.0]*3))
def chunk_ms():
    B, hh, ww = 4, 512, 512
    low = torch.rand(B, 5, hh, ww, device="cuda"); flow = (torch.rand(B, 2, hh, ww, device="cuda") - 0.5) * 0.04
    state = torch.rand(B, 6, 4 * hh, 4 * ww, device="cuda")
    def step():
        _, _, st = h.forward_recurrent(low, state, flow, True, "hybrid", want_raw=False, want_state=True, out_state=state)
        shd(st)
    for _ in range(5): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30): step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 30
def frame_ms():
    lh, lw = 270, 480
    low = torch.rand(1, 5, lh, lw, device="cuda"); flow = (torch.rand(1, 2, lh, lw, device="cuda") - 0.5) * 0.04
    state = torch.rand(1, 6, 4 * lh, 4 * lw, device="cuda"); render = torch.rand(1, 12, lh, lw, device="cuda")
    params = shd.params22()
    def frame():
        _, _, st = h.forward_recurrent(low, state, flow, True, "hybrid", want_raw=False, want_state=True, out_state=state)
        return ops.compose(render, st, params, True, False, 0.0)
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): frame()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): frame()
    for _ in range(20): g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(200):
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(); g.replay(); a1.record(); a1.synchronize(); ts.append(a0.elapsed_time(a1))
    ts.sort()
    return ts[100]
with torch.no_grad():
    for rep in range(3):
        for v in (1, 0):
            isr._lib.set_tunable("pdl_small", v)
            print(f"pdl_small={v} chunk step {chunk_ms():.3f} ms   1080p graph p50 {frame_ms():.4f} ms", flush=True)
isr._lib.set_tunable("pdl_small", 1)
