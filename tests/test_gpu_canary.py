"""Guard-band tests: every output and the network workspace are carved out of a larger sentinel-filled allocation and the
C ABI is called directly on the carved pointers; after the call the bytes on both sides must still hold the sentinel.
(compute-sanitizer is not available on the GPU pool, so out-of-bounds writes are hunted this way, on ragged shapes that
do not divide the kernels' tiles.)"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

PAD = 1 << 16          # guard elements on each side
F_SENT = 1.2345e30


@pytest.fixture(scope="module")
def isr():
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    import isosurfacesuperresolution_b200 as pkg
    return pkg


class Guarded:
    """n elements of ``dtype`` with PAD sentinel elements before and after; ``.t`` is the inner view."""

    def __init__(self, shape, dtype=torch.float32, align=256):
        n = int(np.prod(shape))
        self.sent = F_SENT if dtype == torch.float32 else 0xA5
        self.big = torch.full((n + 2 * PAD + align,), self.sent, dtype=dtype, device="cuda")
        off = PAD
        while (self.big.data_ptr() + off * self.big.element_size()) % align:
            off += 1
        self.off, self.n = off, n
        self.t = self.big[off:off + n].view(shape)

    @property
    def ptr(self):
        return ctypes.c_void_p(self.t.data_ptr())

    def intact(self):
        lo, hi = self.big[:self.off], self.big[self.off + self.n:]
        return bool((lo == self.sent).all()) and bool((hi == self.sent).all())

    def written(self):
        if self.t.dtype == torch.uint8:          # a legitimate byte may equal the sentinel: ask for "almost none"
            return float((self.t == self.sent).float().mean()) < 0.02
        return bool((self.t != self.sent).all())


def P(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def rnd(*shape, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.rand(*shape, device="cuda", generator=g) * 2 - 1


@pytest.mark.parametrize("shape", [(1, 5, 7), (2, 19, 27), (1, 33, 130)])
def test_pointwise_ops_stay_inside_their_outputs(isr, shape):
    from isosurfacesuperresolution_b200._lib import check, lib
    from isosurfacesuperresolution_b200.synthetic import shading_params
    L = lib()
    B, h, w = shape
    H, W = 4 * h, 4 * w
    prm = np.ascontiguousarray(shading_params(), dtype=np.float32)
    pp = prm.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    high = rnd(B, 6, H, W, seed=1)
    flow = rnd(B, 2, h, w, seed=2) * 0.05
    low12 = rnd(B, 12, h, w, seed=3)
    outs = []
    o = Guarded((B, 96, h, w)); outs.append(("flatten_high", o))
    check(L.isr_flatten_high(P(high), o.ptr, B, 6, H, W, 4, stream()))
    o = Guarded((B, 6, H, W)); outs.append(("warp_upscale", o))
    check(L.isr_warp_upscale(P(high), P(flow), o.ptr, B, 6, h, w, 4, 1, 0, stream()))
    o = Guarded((B, 6, H, W)); outs.append(("initial_image", o))
    check(L.isr_initial_image(P(low12[:, :5].contiguous()), o.ptr, B, 5, h, w, 6, 2, 0, 4, stream()))
    o = Guarded((B, 6, H, W)); outs.append(("postprocess", o))
    check(L.isr_postprocess(P(high), o.ptr, B, 6, H, W, 1, stream()))
    o = Guarded((B, 3, H, W)); outs.append(("shade", o))
    check(L.isr_shade(P(high), pp, o.ptr, B, 6, H, W, stream()))
    o = Guarded((B, H, W, 3), torch.uint8); outs.append(("shade_rgb8", o))
    check(L.isr_shade_rgb8(P(high), pp, o.ptr, B, 6, H, W, stream()))
    o = Guarded((B, 12, H, W)); outs.append(("compose", o))
    check(L.isr_compose(P(low12), P(high), pp, o.ptr, B, 12, 6, h, w, 1, 1, ctypes.c_float(0.0), stream()))
    # G-buffer ingest from the interleaved frame: four guarded outputs at once
    inter = low12.permute(0, 2, 3, 1).contiguous()
    g_low, g_flow = Guarded((B, 5, h, w)), Guarded((B, 2, h, w))
    g_hole, g_img = Guarded((B, h, w), torch.uint8), Guarded((B, 12, h, w))
    check(L.isr_ingest_gbuffer(P(inter), 1, B, 12, h, w, 0, g_low.ptr, g_flow.ptr, g_hole.ptr, g_img.ptr, stream()))
    outs += [("ingest.low", g_low), ("ingest.flow", g_flow), ("ingest.hole", g_hole), ("ingest.image12", g_img)]
    # flow in-painting: guarded output AND guarded workspace
    nb = ctypes.c_size_t(0)
    check(L.isr_inpaint_workspace_bytes(B, h, w, ctypes.byref(nb)))
    ws = Guarded((nb.value // 4,))
    o = Guarded((B, 2, h, w)); outs += [("inpaint.out", o), ("inpaint.workspace", ws)]
    hole = (rnd(B, h, w, seed=4) > 0.3).to(torch.uint8)
    check(L.isr_inpaint_flow(P(flow), P(hole), o.ptr, B, h, w, ws.ptr, nb.value, stream()))
    torch.cuda.synchronize()
    for name, g in outs:
        assert g.intact(), name + " wrote outside its buffer"
        if name != "inpaint.workspace":
            assert g.written(), name + " left part of its output untouched"


@pytest.mark.parametrize("shape", [(1, 9, 11), (2, 19, 27), (1, 40, 72)])
@pytest.mark.parametrize("prec", ["hybrid", "bf16x3"])
def test_network_stays_inside_outputs_and_workspace(isr, weights, shape, prec):
    """isr_net_forward / isr_net_forward_recurrent with EXACTLY isr_net_workspace_bytes of workspace, on shapes that do
    not divide the 8 x 16 / 16 x 8 tiles, the 2x / 4x fused-upsample bands or the 16-pixel packing blocks."""
    from argparse import Namespace
    from isosurfacesuperresolution_b200._lib import PREC, check, lib
    L = lib()
    B, h, w = shape
    opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    net = net.cuda().eval()
    with torch.no_grad():
        x = rnd(B, 101, h, w, seed=5)
        ref_recon, ref_raw = net(x)                        # also builds the handle and uploads the weights
    handle = net._handle._net                              # the isr_net* the drop-in module owns
    nb = ctypes.c_size_t(0)
    check(L.isr_net_workspace_bytes(handle, B, h, w, PREC[prec], ctypes.byref(nb)))
    ws = Guarded(((nb.value + 3) // 4,), align=1024)
    recon, raw, state = Guarded((B, 6, 4 * h, 4 * w)), Guarded((B, 6, 4 * h, 4 * w)), Guarded((B, 6, 4 * h, 4 * w))
    check(L.isr_net_forward(handle, P(x), recon.ptr, raw.ptr, state.ptr, 1, B, h, w, PREC[prec], ws.ptr, nb.value, stream()))
    torch.cuda.synchronize()
    for name, g in (("recon", recon), ("raw", raw), ("state", state), ("workspace", ws)):
        assert g.intact(), "isr_net_forward: " + name + " overrun"
    assert recon.written() and raw.written() and state.written()
    if prec == "hybrid":
        assert torch.equal(recon.t, ref_recon) and torch.equal(raw.t, ref_raw)
    # recurrent form
    low, flow = rnd(B, 5, h, w, seed=6), rnd(B, 2, h, w, seed=7) * 0.05
    prev = state.t.clone()
    recon2, state2 = Guarded((B, 6, 4 * h, 4 * w)), Guarded((B, 6, 4 * h, 4 * w))
    ws2 = Guarded(((nb.value + 3) // 4,), align=1024)
    check(L.isr_net_forward_recurrent(handle, P(low), P(prev), P(flow), 5, 6, 1, 0, recon2.ptr, None, state2.ptr, 1, B, h, w,
                                      PREC[prec], ws2.ptr, nb.value, stream()))
    torch.cuda.synchronize()
    for name, g in (("recon", recon2), ("state", state2), ("workspace", ws2)):
        assert g.intact(), "isr_net_forward_recurrent: " + name + " overrun"
    assert recon2.written() and state2.written() and bool(torch.isfinite(recon2.t).all())
    # a workspace that is too small is refused, not overrun
    rc = L.isr_net_forward(handle, P(x), recon.ptr, raw.ptr, None, 1, B, h, w, PREC[prec], ws.ptr, nb.value // 2, stream())
    assert rc != 0


def test_raycaster_stays_inside_its_frame(isr):
    from test_raycast_oracle import make_params, sphere_volume
    vol = torch.from_numpy(sphere_volume(24, [(0, 0, 0)], [0.3])).cuda()
    for (W, H) in ((17, 9), (90, 61), (1, 1)):
        Pm, _, _ = make_params(24, [0.0, 0.1, -1.5], [0, 0, 0], W, H)
        out = Guarded((H, W, 12))
        hemi, rot = isr.ops.raycast_ao_tables()
        for n_ao in (4, 20):                                   # one-pass kernel, then the two-pass form (>= 16 rays)
            out.t.fill_(F_SENT)
            isr.ops.raycast_iso(vol, Pm, out.t, ao_samples=n_ao, hemisphere=hemi, rotations=rot,
                                skip=isr.ops.raycast_build_skip(vol))
            torch.cuda.synchronize()
            assert out.intact() and out.written()


def nan_guarded(t):
    """a copy of ``t`` with NaN on both sides: a read outside the tensor poisons whatever it feeds"""
    big = torch.full((t.numel() + 2 * PAD + 64,), float("nan") if t.dtype.is_floating_point else 0xA5, dtype=t.dtype,
                     device="cuda")
    off = PAD
    while (big.data_ptr() + off * big.element_size()) % 256:
        off += 1
    v = big[off:off + t.numel()].view(t.shape)
    v.copy_(t)
    return v


@pytest.mark.parametrize("shape", [(1, 9, 11), (2, 19, 27), (1, 33, 130)])
def test_no_reads_outside_the_inputs(isr, weights, shape):
    """Inputs surrounded by NaN: every result must equal the one computed from ordinary tensors, bit for bit (a stray read
    of a neighbour would show up as NaN or as a different value)."""
    from argparse import Namespace
    from isosurfacesuperresolution_b200 import ops
    from isosurfacesuperresolution_b200.synthetic import shading_params
    B, h, w = shape
    opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    net = net.cuda().eval()
    low, flow = rnd(B, 5, h, w, seed=11), rnd(B, 2, h, w, seed=12) * 0.2      # large flow: taps leave the image
    prev = rnd(B, 6, 4 * h, 4 * w, seed=13)
    frame = rnd(B, h, w, 12, seed=14)
    prm = shading_params()
    with torch.no_grad():
        a = net(low, prev, flow, special_mask=True, return_state=True)
        b = net(nan_guarded(low), nan_guarded(prev), nan_guarded(flow), special_mask=True, return_state=True)
    for u, v in zip(a, b):
        if u is not None:
            assert torch.isfinite(v).all() and torch.equal(u, v)
    pairs = [
        (ops.warp_upscale(prev, flow, 4, True), ops.warp_upscale(nan_guarded(prev), nan_guarded(flow), 4, True)),
        (ops.flatten_high(prev, 4), ops.flatten_high(nan_guarded(prev), 4)),
        (ops.shade(prev, prm), ops.shade(nan_guarded(prev), prm)),
        (ops.shade_rgb8(prev, prm), ops.shade_rgb8(nan_guarded(prev), prm)),
        (ops.postprocess(prev, True), ops.postprocess(nan_guarded(prev), True)),
        (ops.initial_image(low, 6, "input", False, 4), ops.initial_image(nan_guarded(low), 6, "input", False, 4)),
    ]
    ia, ib = ops.ingest_gbuffer(frame, "unshaded", interleaved=True), ops.ingest_gbuffer(nan_guarded(frame), "unshaded",
                                                                                        interleaved=True)
    pairs += [(x, y) for x, y in zip(ia, ib) if x is not None]
    hole = (rnd(B, h, w, seed=15) > 0.3).to(torch.uint8)
    pairs.append((ops.inpaint_flow(flow, hole), ops.inpaint_flow(nan_guarded(flow), nan_guarded(hole))))
    for i, (u, v) in enumerate(pairs):
        assert torch.equal(u, v), f"pair {i}"
        if v.dtype.is_floating_point:
            assert torch.isfinite(v).all(), f"pair {i}"


def test_calls_follow_the_callers_stream(isr, weights):
    """Every entry point takes the caller's stream (asynchronous, no hidden synchronisation): the same
    recurrence issued on a side stream, interleaved with unrelated work on the default stream, gives the same frames."""
    from argparse import Namespace
    from oracle import oracle as orc
    opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    net = net.cuda().eval()
    low, flow, _ = orc.synthetic_gbuffer(2, 24, 40, 41, T=4)
    low, flow = torch.from_numpy(low).cuda(), torch.from_numpy(flow).cuda()
    runner = isr.SequenceRunner(net, initial_image="input")
    ref, _, ref_state = runner.run(low, flow)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    junk = torch.zeros(4096, 4096, device="cuda")
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        out, _, state = runner.run(low, flow)
    for _ in range(20):                       # the default stream stays busy while the side stream runs the frames
        junk = junk @ junk
    side.synchronize()
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(ref, out)) and torch.equal(ref_state, state)


def test_two_host_threads_with_their_own_networks(isr, weights):
    """The library keeps per-network plans and thread-local error state: two host threads, each with its own module and
    stream, run their recurrences concurrently (ctypes releases the GIL inside the calls) and get the serial results."""
    import threading
    from argparse import Namespace
    from oracle import oracle as orc
    opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
    jobs = []
    for k, (hh, ww) in enumerate(((24, 40), (19, 27))):
        net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
        net.load_state_dict({n: torch.from_numpy(v) for n, v in weights.items()})
        net = net.cuda().eval()
        low, flow, _ = orc.synthetic_gbuffer(2, hh, ww, 50 + k, T=4)
        low, flow = torch.from_numpy(low).cuda(), torch.from_numpy(flow).cuda()
        runner = isr.SequenceRunner(net, initial_image="input")
        ref, _, _ = runner.run(low, flow)
        jobs.append((runner, low, flow, ref))
    torch.cuda.synchronize()
    results, errors = [None, None], []

    def work(i):
        try:
            runner, low, flow, _ = jobs[i]
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.default_stream())
            with torch.cuda.stream(st):
                for _ in range(3):
                    out, _, _ = runner.run(low, flow)
            st.synchronize()
            results[i] = out
        except Exception as e:                                # surfaced in the main thread
            errors.append(e)

    threads = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for i in range(2):
        assert all(torch.equal(a, b) for a, b in zip(jobs[i][3], results[i]))
