"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the committed golden vectors.
Run on the B200 box: python -m pytest tests -m gpu.  Tolerances are the north star's: indexing bit-exact, warp 1e-5,
network outputs 1e-2 max-abs per channel and >= 50 dB PSNR."""
from argparse import Namespace

import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def isr():
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    import isosurfacesuperresolution_b200 as pkg
    return pkg


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def maxabs(a, b):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))))


# ---- indexing: bit-exact ---------------------------------------------------------------------------------------
def test_flatten_high_bit_exact(isr, golden):
    g = golden("flatten_high")
    y = isr.VideoTools.flatten_high(cu(g["x"]), 4).cpu().numpy()
    assert np.array_equal(y, g["y"])
    # full-size property: inverse of pixel_shuffle, 1080p, 6 channels
    x = torch.randn(1, 6, 1080, 1920, device="cuda")
    f = isr.VideoTools.flatten_high(x, 4)
    assert f.shape == (1, 96, 270, 480)
    assert torch.equal(torch.nn.functional.pixel_shuffle(f, 4), x)


# ---- warp: 1e-5 ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("flow", ["zero", "const", "smooth", "wild"])
@pytest.mark.parametrize("special_mask", [False, True])
def test_warp_upscale_golden(isr, golden, flow, special_mask):
    g = golden("warp_upscale")
    out = isr.VideoTools.warp_upscale(cu(g["img"]), cu(g[f"flow_{flow}"]), 4, special_mask=special_mask)
    assert maxabs(out, g[f"{flow}_{int(special_mask)}"]) <= 1e-5


def test_warp_upscale_rgb_nonsquare(isr, golden):
    g = golden("warp_upscale_rgb")
    assert maxabs(isr.VideoTools.warp_upscale(cu(g["img"]), cu(g["flow"]), 4), g["out"]) <= 1e-5


def test_warp_upscale_1080p_vs_oracle(isr):
    low, flow, ao = orc.synthetic_gbuffer(1, 270, 480, 33)
    rng = np.random.RandomState(3)
    img = rng.uniform(-1, 1, (1, 6, 1080, 1920)).astype(np.float32)
    ref = orc.warp_upscale(img, flow[:, 0], 4, special_mask=True, backend="c")
    out = isr.VideoTools.warp_upscale(cu(img), cu(flow[:, 0]), 4, special_mask=True)
    assert maxabs(out, ref) <= 1e-5


def test_warp_align_corners_identity(isr):
    # with the as-trained align_corners=True convention a zero flow is the identity warp
    img = torch.rand(2, 3, 64, 96, device="cuda")
    out = isr.VideoTools.warp_upscale(img, torch.zeros(2, 2, 16, 24, device="cuda"), 4, align_corners=True)
    assert maxabs(out, img) <= 2e-5


# ---- initial image / post-process / shading --------------------------------------------------------------------
def test_initial_image(isr, golden):
    g = golden("initial_image")
    cur = cu(g["cur"])
    for mode in ("zero", "unshaded", "input"):
        for ch in (5, 6):
            assert maxabs(isr.initialImage(cur, ch, mode, False, 4), g[f"{mode}_{ch}"]) <= 1e-6
    assert maxabs(isr.initialImage(cur, 6, "unshaded", True, 4), g["unshaded_6_inv"]) == 0
    assert maxabs(isr.initialImage(cur, 3, "input", False, 4), g["input_3"]) <= 1e-6
    with pytest.raises(ValueError):
        isr.initialImage(cur, 4, "unshaded", False, 4)
    with pytest.raises(ValueError):
        isr.initialImage(cur, 6, "bogus", False, 4)


def _shading_from_params(isr, p):
    s = isr.ScreenSpaceShading(torch.device("cuda"))
    s.fov(30)
    s.ambient_light_color(np.array(p[0:3]))
    s.diffuse_light_color(np.array(p[3:6]))
    s.specular_light_color(np.array(p[6:9]))
    s.specular_exponent(int(p[19]))
    s.light_direction(np.array(p[9:12], dtype=np.float64))
    s.material_color(np.array(p[12:15]))
    s.background(np.array(p[15:18]))
    s.ambient_occlusion(float(p[18]))
    s.enable_specular = bool(p[20])
    s.inverse_ao = bool(p[21])
    return s


@pytest.mark.parametrize("cfg", ["train", "gui", "nospec"])
def test_shading_golden(isr, golden, cfg):
    g = golden("shading")
    s = _shading_from_params(isr, g[f"{cfg}_params"])
    gb = cu(g["gbuffer"])
    assert maxabs(s(gb), g[f"{cfg}_c6"]) <= 1e-5
    assert maxabs(s(gb[:, 0:5].contiguous()), g[f"{cfg}_c5"]) <= 1e-5


def test_shading_1080p_vs_oracle(isr):
    low, _, ao = orc.synthetic_gbuffer(1, 270, 480, 12)
    gb = np.concatenate([low[:, 0], ao[:, 0]], axis=1)
    gb = np.repeat(np.repeat(gb, 4, axis=2), 4, axis=3) + np.random.RandomState(1).normal(0, 0.1, (1, 6, 1080, 1920))
    prm = orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5, background=(1.0, 1.0, 1.0))
    s = _shading_from_params(isr, prm)
    assert maxabs(s(cu(gb)), orc.shade(gb, prm)) <= 1e-5


def test_postprocess(isr, golden):
    from isosurfacesuperresolution_b200 import ops
    g = golden("shading")
    assert maxabs(ops.postprocess(cu(g["gbuffer"]), True), g["postprocess"]) <= 1e-6
    x = cu(g["gbuffer"])
    assert maxabs(ops.postprocess(x, False), torch.clamp(x, 0, 1)) == 0


# ---- convolution kernel ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(1, 16, 8), (2, 37, 29), (1, 270, 480)])
@pytest.mark.parametrize("mode", [("bf16", 2, 2, 2e-4), ("fp16", 2, 2, 5e-5), ("fp16", 1, 1, 2e-2), ("bf16", 1, 1, 1.5e-1)])
def test_conv3x3_vs_fp64(isr, shape, mode):
    from isosurfacesuperresolution_b200 import ops
    fmt, na, nw, tol = mode
    B, H, W = shape
    g = torch.Generator().manual_seed(11)
    x = (torch.rand(B, 64, H, W, generator=g) * 2 - 1).cuda()
    w = (torch.randn(64, 64, 3, 3, generator=g) * 0.06).cuda()
    b = torch.randn(64, generator=g).cuda()
    ref = torch.nn.functional.conv2d(x.double(), w.double(), b.double(), padding=1)
    out = ops.debug_conv3x3(x, w, b, fmt=fmt, n_act_parts=na, n_w_parts=nw)
    assert maxabs(out, ref) <= tol
    add = torch.randn(B, 64, H, W, generator=g).cuda()
    out = ops.debug_conv3x3(x, w, b, addend=add, relu=True, fmt=fmt, n_act_parts=na, n_w_parts=nw)
    assert maxabs(out, torch.relu(ref + add)) <= tol


# ---- network -----------------------------------------------------------------------------------------------------
def _net(isr, weights, cin=101, cout=6, mask=(0, 1, 2, 3, 4), ups="bilinear", recon="residual"):
    opt = Namespace(upsample=ups, reconType=recon, useBN=False)
    net = isr.createNetwork("EnhanceNet", 4, cin, list(mask), cout, opt)
    if weights is not None:
        net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    return net.cuda().eval()


def _check_net(out, ref, tol=1e-2, min_psnr=50.0):
    out = out.detach().cpu().numpy()
    for c in range(ref.shape[1]):
        assert float(np.max(np.abs(out[:, c] - ref[:, c]))) <= tol, f"channel {c}"
    assert orc.psnr(out, ref) >= min_psnr


@pytest.mark.parametrize("ups", ["bilinear", "nearest"])
@pytest.mark.parametrize("prec", ["hybrid", "mixed", "mixed3", "mixed4", "bf16x3", "fp16x3"])
def test_enhancenet_forward_golden(isr, golden, weights, ups, prec):
    g = golden("enhancenet_forward")
    net = _net(isr, weights, ups=ups)
    net.precision = prec
    with torch.no_grad():
        recon, raw = net(cu(g["x"]))
    _check_net(raw, g[f"raw_{ups}"])
    _check_net(recon, g[f"recon_{ups}"])


def test_enhancenet_state_dict_roundtrip(isr, weights):
    net = _net(isr, weights)
    sd = net.state_dict()
    assert list(sd.keys()) == orc.STATE_DICT_KEYS
    for k in sd:
        assert np.array_equal(sd[k].cpu().numpy(), weights[k])
    first = net
    while list(first.children()):
        first = list(first.children())[0]
    assert first.in_channels == 101                    # LoadedModel's layout sniffing (loadedmodel.py:26-35)


def test_enhancenet_direct_and_shaded(isr, golden, weights):
    g = golden("enhancenet_forward")
    net = _net(isr, weights, recon="direct")
    with torch.no_grad():
        recon, raw = net(cu(g["x"]))
    _check_net(recon, g["recon_direct"])
    assert torch.equal(recon, raw)
    # shaded variant (56 -> 3): weights rebuilt by the same seeded constructor as the reference's
    torch.manual_seed(1234)
    net_s = _net(isr, None, cin=56, cout=3, mask=(0, 1, 2))
    ws = [float(sum(v.double().sum() for v in net_s.state_dict().values())),
          float(sum(v.double().abs().sum() for v in net_s.state_dict().values()))]
    assert np.allclose(ws, g["shaded_weight_sum"], rtol=1e-6)
    with torch.no_grad():
        rs, raws = net_s(cu(g["shaded_x"]))
    _check_net(raws, g["shaded_raw"])
    _check_net(rs, g["shaded_recon"])


def test_enhancenet_c1_vs_oracle(isr, weights):
    """Config 1: 128x128 -> 512x512, uniform stress input and synthetic G-buffer input, vs the fp32 CPU oracle."""
    net = _net(isr, weights)
    rng = np.random.RandomState(7)
    x = rng.uniform(-1, 1, (1, 101, 128, 128)).astype(np.float32)
    ref_recon, ref_raw = orc.enhancenet_forward(weights, x, "bilinear", "residual", 5, backend="torch")
    with torch.no_grad():
        recon, raw = net(cu(x))
    _check_net(raw, ref_raw)
    _check_net(recon, ref_recon)


def test_sequence_recurrence_golden(isr, golden, weights):
    """3-frame recurrence against the live-reference fixture.  Per-frame parity (north star: 1e-2 / 50 dB given the same
    inputs) is checked TEACHER-FORCED: frame j is computed from the reference's own previous output.  Free-running, the
    feedback loop amplifies any operand-rounding difference about 2x per frame with these random-init weights (measured:
    profiles/r01_precision_sweep.json), so the free-running check uses the split-precision mode everywhere (bf16x3) at
    the same 1e-2, and the default mode at a documented looser bound."""
    g = golden("sequence_unshaded")
    net = _net(isr, weights)
    shd = _shading_from_params(isr, g["shade_params"])
    low, flow = cu(g["low"]), cu(g["flow"])
    # teacher-forced, default precision
    with torch.no_grad():
        for j in range(3):
            if j == 0:
                prev = isr.initialImage(low[:, 0].contiguous(), 6, "input", False, 4)
                pred, _, state = net(low[:, 0].contiguous(), prev, None, return_state=True)
            else:
                ref_state = cu(orc.postprocess(g[f"pred_{j - 1}"], unshaded=True))
                pred, _, state = net(low[:, j].contiguous(), ref_state, flow[:, j - 1].contiguous(), special_mask=True,
                                     return_state=True)
            _check_net(pred, g[f"pred_{j}"])
            assert maxabs(shd(state), g[f"rgb_{j}"]) <= 2e-2
    # free-running: see test_gpu_parity_r02.py::test_sequence_unshaded_10_frames_free_running for the bound -- the split
    # scheme stays inside the fixed bar, the cheaper schemes inside k x (their own single-frame error) x (the measured
    # amplification of the oracle's own feedback loop); no hard-coded looser tolerance
    from test_gpu_parity_r02 import _amplification
    ref = [g[f"pred_{j}"] for j in range(3)]
    gain = _amplification(weights, g["low"], g["flow"], ref)
    for prec in ("bf16x3", "mixed", "hybrid"):
        net.precision = prec
        runner = isr.SequenceRunner(net, initial_image="input", shading=shd)
        preds, rgbs, _ = runner.run(low, flow)
        e1 = maxabs(preds[0], ref[0])
        assert e1 <= 1e-2
        for j in range(3):
            bound = 1e-2 if prec == "bf16x3" else 3.0 * max(gain[j] / gain[0], 1.0) * e1
            assert maxabs(preds[j], ref[j]) <= bound, (prec, j)
            assert maxabs(rgbs[j], g[f"rgb_{j}"]) <= 2 * bound


def test_fused_recurrent_equals_unfused(isr, weights):
    """forward(low, prev, flow) == forward(cat(low, flatten_high(warp_upscale(prev, flow)))) -- same kernels downstream,
    so the packed inputs must agree to the last bit of the 16-bit operands."""
    net = _net(isr, weights)
    low, flow, ao = orc.synthetic_gbuffer(2, 24, 40, 5, T=2)
    prev = torch.rand(2, 6, 96, 160, device="cuda") * 2 - 1
    with torch.no_grad():
        a, _ = net(cu(low[:, 1]), prev, cu(flow[:, 0]), special_mask=True)
        warped = isr.VideoTools.warp_upscale(prev, cu(flow[:, 0]), 4, special_mask=True)
        single = torch.cat((cu(low[:, 1]), isr.VideoTools.flatten_high(warped, 4)), dim=1)
        b, _ = net(single)
    assert torch.equal(a, b)


def test_sequence_runner_chunks_are_bit_identical(isr, weights):
    """configs[4]-style micro-batching: chunks of sequences give exactly the unchunked result."""
    net = _net(isr, weights)
    low, flow, _ = orc.synthetic_gbuffer(5, 24, 32, 31, T=3)
    runner = isr.SequenceRunner(net, initial_image="input", shading=_shading_from_params(isr, orc.shading_params()))
    p_all, r_all, s_all = runner.run(cu(low), cu(flow))
    p_ch, r_ch, s_ch = runner.run(cu(low), cu(flow), max_batch=2)
    assert len(p_ch) == 3 and len(r_ch) == 3 and p_ch[0].shape[0] == 5
    assert all(torch.equal(a, b) for a, b in zip(p_all, p_ch)) and all(torch.equal(a, b) for a, b in zip(r_all, r_ch))
    assert torch.equal(s_all, s_ch)


def test_output_conv_single_weight_plane_tunable(isr, golden, weights):
    """final_w_parts = 1 drops the low weight plane of the output conv (-1.7 % step time, +0.4e-3 error,
    profiles/r01_precision_sweep_final_w_parts.json); not the default, but it must stay inside the parity bar.  Tunables
    are per handle (isr_net_set_tunable)."""
    g = golden("enhancenet_forward")
    net = _net(isr, weights)
    old = net._native().set_tunable("final_w_parts", 1)
    with torch.no_grad():
        recon, _ = net(cu(g["x"]))
    _check_net(recon, g["recon_bilinear"])
    net._native().set_tunable("final_w_parts", old)
    with torch.no_grad():
        recon2, _ = net(cu(g["x"]))
    assert not torch.equal(recon, recon2)          # the tunable did change the arithmetic, and was restored
    _check_net(recon2, g["recon_bilinear"])


def test_loaded_model_facade(isr, weights):
    net = _net(isr, weights)
    lm = isr.LoadedModel({"model": net, "parameters": {"initialImage": "input"}}, torch.device("cuda"), 4)
    assert lm.unshaded and lm.input_channels == 101 and lm.initial_image_mode == "input"
    low, flow, ao = orc.synthetic_gbuffer(1, 32, 48, 8, T=2)
    frame = np.zeros((1, 12, 32, 48), np.float32)
    frame[:, 3] = low[:, 0, 0] * 0.5 + 0.5
    frame[:, 4:8] = low[:, 0, 1:5]
    frame[:, 8:10] = flow[:, 0]
    p0 = lm.inference(cu(frame), None)
    p1 = lm.inference(cu(frame), p0)
    assert p0.shape == (1, 6, 128, 192) and p1.shape == p0.shape and torch.isfinite(p1).all()
    # frame 0 against the oracle
    inp = np.concatenate([frame[:, 3:4] * 2 - 1, frame[:, 4:8]], axis=1)
    prev = orc.initial_image(inp, 6, "input", False, 4)
    single = np.concatenate([inp, orc.flatten_high(prev, 4)], axis=1)
    ref, _ = orc.enhancenet_forward(weights, single, "bilinear", "residual", 5, backend="torch")
    _check_net(p0, ref)


def test_no_cpu_fallback(isr, weights):
    net = _net(isr, weights).cpu()
    with torch.no_grad(), pytest.raises(RuntimeError):
        net(torch.zeros(1, 101, 8, 8))
    with pytest.raises(RuntimeError):
        isr.VideoTools.flatten_high(torch.zeros(1, 6, 8, 8), 4)


@pytest.mark.parametrize("prec", ["hybrid", "mixed", "bf16x3"])
def test_output_conv_tap_stacked_equals_generic(isr, golden, weights, prec):
    """The tap-stacked output-conv kernel (isr_conv_final.cuh) and the generic N=16 kernel compute the same convolution
    (different fp32 summation order only); same for the N=128 hi/lo weight stacking of the split-precision layers."""
    from isosurfacesuperresolution_b200 import _lib
    g = golden("enhancenet_forward")
    x = cu(g["x"])
    outs = {}
    try:
        for fin, stk in ((3, 1), (2, 1), (2, 0), (1, 1), (0, 1), (1, 0), (0, 0)):
            _lib.set_tunable("final_stacked", fin)
            _lib.set_tunable("stack_w", stk)
            net = _net(isr, weights)
            net.precision = prec
            with torch.no_grad():
                outs[(fin, stk)] = net(x)[0].clone()
    finally:
        _lib.set_tunable("final_stacked", 3)
        _lib.set_tunable("stack_w", 1)
    # column-stacked output conv with the register gather (isr_conv_final2.cuh, the default) vs the generic kernel
    assert maxabs(outs[(3, 1)], outs[(0, 1)]) <= 5e-5
    # tap-stacked vs generic output conv: same products, different fp32 summation order
    assert maxabs(outs[(1, 1)], outs[(0, 1)]) <= 5e-5 and maxabs(outs[(1, 0)], outs[(0, 0)]) <= 5e-5
    # CTA-pair output conv (tunable final_stacked=2) vs the generic single-CTA kernel
    assert maxabs(outs[(2, 1)], outs[(0, 1)]) <= 5e-5 and maxabs(outs[(2, 0)], outs[(0, 0)]) <= 5e-5
    # hi/lo weight stacking moves the small hi*lo products into their own accumulator columns; the tensor core's fp32
    # accumulation is not order-independent at the 1e-4 level (measured 1.2e-4 in bf16x3, the stacked form being the
    # closer one to the fp32 reference), and in `mixed` the fp16 tail re-quantises that difference to its own noise level
    assert maxabs(outs[(1, 1)], outs[(1, 0)]) <= (5e-4 if prec == "bf16x3" else 6e-3)
    for k, v in outs.items():
        _check_net(v, g["recon_bilinear"])


@pytest.mark.parametrize("prec", ["hybrid", "mixed", "bf16x3", "fp16"])
def test_cta_pair_kernel_equals_single_cta(isr, golden, weights, prec):
    """conv3x3_pair_kernel (tcgen05 cta_group::2, M=256 over two CTAs) computes the same GEMM as the single-CTA kernel:
    same products, same accumulation order per output element."""
    from isosurfacesuperresolution_b200 import _lib
    g = golden("enhancenet_forward")
    x = cu(g["x"])
    outs = {}
    try:
        _lib.set_tunable("fuse_up", 0)          # like for like: the fused upsample + conv exists in pair form only
        for pair in (1, 0):
            _lib.set_tunable("cta_pair", pair)
            net = _net(isr, weights)
            net.precision = prec
            with torch.no_grad():
                outs[pair] = net(x)[0].clone()
    finally:
        _lib.set_tunable("cta_pair", 1)
        _lib.set_tunable("fuse_up", 1)
    assert maxabs(outs[1], outs[0]) <= (6e-3 if prec in ("mixed", "hybrid") else 1e-4)
    if prec != "fp16":
        _check_net(outs[1], g["recon_bilinear"])


def test_host_pipeline_equals_device_recurrence(isr, weights):
    """isr_pipe_* (pinned host buffers, three streams, depth-2 staging) produces exactly the frames of the device-resident
    recurrence: same kernels, only the copies are overlapped."""
    from isosurfacesuperresolution_b200 import ops
    net = _net(isr, weights)
    B, T, h, w = 2, 5, 24, 40
    low, flow, _ = orc.synthetic_gbuffer(B, h, w, 21, T=T)
    preds, _, _ = isr.SequenceRunner(net, initial_image="input").run(cu(low), cu(flow))
    pipe = ops.RecurrentPipeline(net._native(), 5, 6, True, "input", B, h, w, net.precision, depth=2)
    outs = [torch.empty((B, 6, 4 * h, 4 * w)).pin_memory() for _ in range(T)]
    lows = [torch.from_numpy(np.ascontiguousarray(low[:, j])).pin_memory() for j in range(T)]
    flows = [torch.from_numpy(np.ascontiguousarray(flow[:, j])).pin_memory() for j in range(T)]
    for j in range(T):
        pipe.submit(lows[j], None if j == 0 else flows[j - 1], outs[j])
    pipe.wait()
    for j in range(T):
        assert torch.equal(outs[j], preds[j].cpu()), f"frame {j}"
    with pytest.raises(ValueError):
        pipe.submit(lows[0][:, :4].contiguous(), None, outs[0])
    pipe.close()


@pytest.mark.parametrize("masking", [False, True])
def test_compose_frame_vs_oracle(isr, masking):
    """a13: GUI composition (12-channel bilinear x4 + shading + overwrite + mask lerp) in one kernel."""
    rng = np.random.RandomState(3)
    low, flow, ao = orc.synthetic_gbuffer(1, 54, 96, 8)
    img = rng.uniform(0, 1, (1, 12, 54, 96)).astype(np.float32)
    img[:, 3:8] = low[:, 0]
    raw = orc.postprocess(rng.uniform(-1.2, 1.2, (1, 6, 216, 384)).astype(np.float32), unshaded=True)
    prm = orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5, background=(0.2, 0.3, 0.4))
    shd = _shading_from_params(isr, prm)
    ref = orc.compose_frame(img, raw, prm, masking=masking, background0=0.2)
    out = isr.compose_frame(cu(img), cu(raw), shd, masking=masking, background=(0.2, 0.3, 0.4))
    assert out.shape == (1, 12, 216, 384)
    assert maxabs(out, ref) <= 1e-5
    # shaded networks: RGB prediction clamped into channels 0:3, everything else upsampled
    rgb = rng.uniform(-0.2, 1.2, (1, 3, 216, 384)).astype(np.float32)
    ref_s = orc.compose_frame(img, rgb, None, masking=masking, background0=1.0)
    out_s = isr.compose_frame(cu(img), cu(rgb), None, masking=masking, background=(1.0, 1.0, 1.0))
    assert maxabs(out_s, ref_s) <= 1e-5


# ---- in front of the network: ingest + device in-painting -------------------------------
def _render12(B=2, h=45, w=70, seed=3):
    rng = np.random.RandomState(seed)
    r = rng.uniform(-0.2, 1.2, (B, h, w, 12)).astype(np.float32)
    r[..., 3] = (rng.rand(B, h, w) > 0.4).astype(np.float32)
    r[..., 8:10] = rng.uniform(-0.02, 0.02, (B, h, w, 2)).astype(np.float32)
    return r


@pytest.mark.parametrize("variant", ["unshaded", "shaded", "shaded_depth", "shaded_normal", "shaded_normal_depth"])
def test_ingest_gbuffer_bit_exact(isr, variant):
    from isosurfacesuperresolution_b200 import ops
    r = _render12()
    v = ops.INGEST_VARIANTS[variant]
    ref = orc.ingest_gbuffer(r, v, interleaved=True)
    for interleaved, src in ((True, r), (False, np.ascontiguousarray(r.transpose(0, 3, 1, 2)))):
        low, flow, hole, img = ops.ingest_gbuffer(cu(src), variant, interleaved=interleaved, want_image12=True)
        assert np.array_equal(low.cpu().numpy(), ref[0])               # slicing / remap / clamp: bit-exact
        assert np.array_equal(flow.cpu().numpy(), ref[1])
        assert np.array_equal(hole.cpu().numpy(), ref[2])
        assert np.array_equal(img.cpu().numpy(), ref[3])
    # 10-channel planar render (no AO / shadow): image12 is refused, the rest works
    low, flow, hole, _ = ops.ingest_gbuffer(cu(np.ascontiguousarray(r.transpose(0, 3, 1, 2)[:, :10])), variant, interleaved=False)
    assert np.array_equal(low.cpu().numpy(), ref[0])
    with pytest.raises(Exception):
        ops.ingest_gbuffer(cu(np.ascontiguousarray(r.transpose(0, 3, 1, 2)[:, :10])), variant, interleaved=False, want_image12=True)


@pytest.mark.parametrize("shape", [(1, 37, 53), (2, 64, 96), (1, 1, 7), (1, 1, 1), (1, 270, 480)])
def test_inpaint_flow_vs_oracle(isr, shape):
    from isosurfacesuperresolution_b200 import ops
    B, h, w = shape
    rng = np.random.RandomState(h * 1000 + w)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    flow = np.stack([0.01 * np.sin(xx / 9) + 0.002 * yy / max(h, 1), 0.01 * np.cos(yy / 7)], 0)[None].repeat(B, 0)
    flow = (flow + rng.uniform(-1e-3, 1e-3, flow.shape)).astype(np.float32)
    hole = (rng.rand(B, h, w) < 0.35).astype(np.uint8)
    if h > 30:
        hole[:, 10:25, 20:40] = 1                                      # a large hole: filled from coarse levels
    ref = orc.inpaint_flow_pushpull(flow, hole)
    out = ops.inpaint_flow(cu(flow), torch.from_numpy(hole).cuda()).cpu().numpy()
    assert np.array_equal(out[:, :, :, :][np.broadcast_to((hole == 0)[:, None], out.shape)],
                          flow[np.broadcast_to((hole == 0)[:, None], flow.shape)])     # valid pixels untouched
    assert maxabs(out, ref) <= 1e-7                                    # same fp32 operations in the same order
    # degenerate masks
    assert torch.equal(ops.inpaint_flow(cu(flow), torch.zeros(B, h, w, dtype=torch.uint8, device="cuda")), cu(flow))
    assert not ops.inpaint_flow(cu(flow), torch.ones(B, h, w, dtype=torch.uint8, device="cuda")).any()


def test_loaded_model_gpu_inpaint_and_interleaved(isr, weights):
    """LoadedModel.inference with the device in-painter and with the renderer's interleaved frame: same result as the
    explicit ingest -> in-paint -> network pipeline, nothing leaves the GPU."""
    from isosurfacesuperresolution_b200 import ops
    net = _net(isr, weights)
    lm = isr.LoadedModel({"model": net, "parameters": {"initialImage": "input"}}, torch.device("cuda"), 4)
    low, flow, ao = orc.synthetic_gbuffer(1, 32, 48, 8, T=2)
    frame = np.zeros((1, 12, 32, 48), np.float32)
    frame[:, 3] = low[:, 0, 0] * 0.5 + 0.5
    frame[:, 4:8] = low[:, 0, 1:5]
    frame[:, 8:10] = flow[:, 0] * (frame[:, 3:4] != 0)                 # the renderer writes no flow outside the surface
    p0 = lm.inference(cu(frame), None, inpaint="gpu")
    p1 = lm.inference(cu(frame), p0, inpaint="gpu")
    hwc = np.ascontiguousarray(frame.transpose(0, 2, 3, 1))
    q1 = lm.inference(cu(hwc), p0, inpaint="gpu")                      # interleaved wire format
    assert torch.equal(p1, q1)
    inp, fl, hole, _ = ops.ingest_gbuffer(cu(frame), "unshaded", interleaved=False)
    with torch.no_grad():
        ref1, _ = net(inp, p0, ops.inpaint_flow(fl, hole), special_mask=True)
    assert torch.equal(p1, ref1)
    # the OpenCV path (reference behaviour) stays available; the two fills differ inside the holes (different algorithms,
    # parity unpinned), which the random-init network amplifies at silhouette pixels: compare on average only
    o1 = lm.inference(cu(frame), p0, inpaint="opencv")
    assert torch.isfinite(o1).all() and float((o1 - p1).abs().mean()) <= 0.15
    cvf = lm._inpaint_flow(cu(frame))
    assert maxabs(cvf, ops.inpaint_flow(fl, hole)) <= float(fl.max() - fl.min())      # both fills stay inside the flow range
    with pytest.raises(ValueError):
        lm.inference(cu(frame), p0, inpaint="bogus")


def test_dataset_evaluate_on_npy_sequences(isr, weights, tmp_path):
    """sequences in the reference's .npy format through the fused recurrent path.  `high` holds the CPU
    oracle's own (post-processed) predictions, so every PSNR must be at the operand-rounding noise level (>= 45 dB free-running)."""
    import os
    net = _net(isr, weights)
    for i, seed in enumerate((21, 22, 23)):
        low, flow, _ = orc.synthetic_gbuffer(1, 24, 32, seed, T=3)
        preds = orc.unshaded_sequence(weights, low, flow, "bilinear", "input", backend="torch")[0]
        # frame j is warped from frame j-1 with flow[:, j-1] (mainVideoUnshaded.py:429-433), as SequenceRunner does
        np.save(os.path.join(tmp_path, "low_%05d.npy" % i), low[0])
        np.save(os.path.join(tmp_path, "flow_%05d.npy" % i), flow[0])
        np.save(os.path.join(tmp_path, "high_%05d.npy" % i),
                np.concatenate([orc.postprocess(p, unshaded=True) for p in preds], axis=0))
    ds = isr.SequenceDataset(str(tmp_path))
    assert len(ds) == 3 and ds.num_frames == 3 and ds.upscale_factor == 4
    res = isr.evaluate(net, ds, batch_size=2, initial_image="input")
    assert res["frames"] == 9
    for k in ("psnr_mask", "psnr_normal", "psnr_depth", "psnr_ao"):
        assert res[k] >= 45.0, (k, res)


# ---- screen-space ambient occlusion ----------------------------------------------------------
@pytest.mark.parametrize("cfg", [dict(), dict(samples=32, sample_radius=0.05), dict(noise_samples=2, bias=0.0)])
def test_ssao_equals_oracle(isr, cfg):
    """isr_ssao against the CPU restatement of ShadowFX::screenSpaceAmbientOcclusion (bit-exact: same fp32 / fp64
    operations in the same order) and, where it was built, against the reference's own ShadowFX.cpp (oracle/_ref)."""
    from isosurfacesuperresolution_b200 import ops
    from test_ssao_oracle import ssao_frame
    for shape in ((48, 64), (67, 45)):
        f = ssao_frame(*shape)
        ref = orc.ssao(f, **cfg)
        out = ops.ssao(cu(f), **cfg).cpu().numpy()
        assert np.array_equal(out[..., :10], f[..., :10]) and np.array_equal(out[..., 11], f[..., 11])   # in place, AO only
        assert np.array_equal(out[..., 10], ref[..., 10])
        if orc.ssao_ref_available():
            assert np.array_equal(out, orc.ssao(f, use_reference_build=True, **cfg))
    # batched interleaved frames and argument checks
    fb = np.stack([ssao_frame(32, 40, seed=s) for s in (1, 2)])
    ob = ops.ssao(cu(fb)).cpu().numpy()
    for i in range(2):
        assert np.array_equal(ob[i], orc.ssao(fb[i]))
    with pytest.raises(Exception):
        ops.ssao(cu(fb), samples=129)


# ---- full BASELINE.json sizes through a size-independent property ---------------------------------------------------
@pytest.mark.parametrize("cfg", ["C2", "C3", "C4", "C5"])
def test_full_size_crop_parity(isr, weights, cfg):
    """The network is translation-covariant with a finite receptive field (21 low-res convs + the three high-res ones and
    the bilinear steps: < 28 low-res pixels), so far enough from a crop's border the output of the FULL-size problem must
    equal the fp32 CPU oracle run on the crop alone.  This carries the 1e-2 / 50 dB bar to the sizes of
    BASELINE.json configs[1] (16 x 256^2), configs[2] (1 x 270x480, ragged tiles) and configs[3] (shaded 56 -> 3,
    4 x 540x960 -> 2160x3840), where the oracle itself would take minutes.  Also: a sequence's result does not depend on
    the rest of its batch (bit-exact)."""
    M, S = 28, 84                                                    # margin, crop side (low-res)
    if cfg == "C4":
        torch.manual_seed(1234)
        net = _net(isr, None, cin=56, cout=3, mask=(0, 1, 2))
        wts = {k: v.detach().cpu().numpy() for k, v in net.state_dict().items()}
        B, h, w, cin, nm = 4, 540, 960, 56, 3
    else:
        net, wts, cin, nm = _net(isr, weights), weights, 101, 5
        # C5: one 4-sequence chunk of configs[4] (64 sequences of 512^2 -> 2048^2 per GPU, SequenceRunner max_batch)
        B, h, w = {"C2": (16, 256, 256), "C3": (1, 270, 480), "C5": (4, 512, 512)}[cfg]
    g = torch.Generator(device="cuda").manual_seed(77)
    x = torch.rand(B, cin, h, w, device="cuda", generator=g) * 2 - 1
    with torch.no_grad():
        recon, raw = net(x)
    assert recon.shape == (B, nm if cfg == "C4" else 6, 4 * h, 4 * w)
    for b, y0, x0 in ((B - 1, h - S, w - S), (0, 0, 0), (B // 2, (h - S) // 2 + 3, (w - S) // 2 + 5)):
        crop = x[b:b + 1, :, y0:y0 + S, x0:x0 + S].cpu().numpy()
        ref_recon, ref_raw = orc.enhancenet_forward(wts, crop, "bilinear", "residual", nm, backend="torch")
        # interior of the crop, except where the crop border IS the image border (same zero padding on both sides)
        ya, yb = (0 if y0 == 0 else M), (S if y0 + S == h else S - M)
        xa, xb = (0 if x0 == 0 else M), (S if x0 + S == w else S - M)
        sl = (slice(None), slice(None), slice(4 * ya, 4 * yb), slice(4 * xa, 4 * xb))
        full = (slice(b, b + 1), slice(None), slice(4 * (y0 + ya), 4 * (y0 + yb)), slice(4 * (x0 + xa), 4 * (x0 + xb)))
        _check_net(raw[full], ref_raw[sl])
        _check_net(recon[full], ref_recon[sl])
    if B > 1:                                                       # batch independence, bit for bit
        with torch.no_grad():
            alone, _ = net(x[B - 1:B].contiguous())
        assert torch.equal(alone, recon[B - 1:B])


@pytest.mark.parametrize("shape", [(1, 32, 48), (2, 19, 27), (1, 270, 480)])
def test_fused_upsample_conv_equals_unfused(isr, weights, shape):
    """conv3x3(bilinear_up2(x)) as four phase convolutions on the low-res tensor with composite weights + the four
    border bands (isr_up2.cu, tunable fuse_up=1, default) against the materialised upsample + conv (fuse_up=0): same
    layer up to operand rounding (the composite weights are rounded once, the unfused path rounds the upsampled
    activations), everywhere including the 2-px border ring and ragged tiles; both meet the oracle bar."""
    from isosurfacesuperresolution_b200 import _lib
    B, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(B, 101, h, w, device="cuda", generator=g) * 2 - 1
    outs = {}
    try:
        for fuse in (1, 0):
            _lib.set_tunable("fuse_up", fuse)
            net = _net(isr, weights)
            with torch.no_grad():
                outs[fuse] = net(x)[0].clone()
    finally:
        _lib.set_tunable("fuse_up", 1)
    d = (outs[1] - outs[0]).abs()
    assert float(d.max()) <= 6e-3
    ring = torch.ones_like(d, dtype=torch.bool)
    ring[:, :, 8:-8, 8:-8] = False
    # the border is not worse than the interior (a wrong ring would be off by O(0.1))
    assert float(d[ring].max()) <= 6e-3 and float(d[ring].mean()) <= 2 * float(d[~ring].mean()) + 1e-5
    if h * w <= 2000:
        ref, _ = orc.enhancenet_forward(weights, x.cpu().numpy(), "bilinear", "residual", 5, backend="torch")
        _check_net(outs[1], ref)
        _check_net(outs[0], ref)


def test_display_pipeline_rgb8(isr, weights):
    """Display form of the host pipeline: the 8-bit shaded frames that reach the host equal
    uint8(clip(ScreenSpaceShading(state) * 255, 0, 255)) (mainGUI.py:507 / mainComparisonVideo3.py:589) of the
    device-resident recurrence, frame for frame; isr_shade_rgb8 equals the fp32 shading + the host conversion."""
    from isosurfacesuperresolution_b200 import ops
    net = _net(isr, weights)
    B, T, h, w = 2, 4, 24, 40
    low, flow, _ = orc.synthetic_gbuffer(B, h, w, 22, T=T)
    params = orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5)
    shd = _shading_from_params(isr, params)
    # reference frames: device recurrence, fp32 shading, host-side conversion as in the reference
    state, refs = None, []
    with torch.no_grad():
        for j in range(T):
            inp = cu(low[:, j])
            if j == 0:
                prev = isr.initialImage(inp, 6, "input", False, 4)
                _, _, state = net(inp, prev, None, return_state=True)
            else:
                _, _, state = net(inp, state, cu(flow[:, j - 1]), special_mask=True, return_state=True)
            rgb = shd(state).cpu().numpy()
            refs.append(np.clip(rgb * 255, 0, 255).astype(np.uint8).transpose(0, 2, 3, 1))
            assert np.array_equal(ops.shade_rgb8(state, shd.params22()).cpu().numpy(), refs[-1])
    pipe = ops.RecurrentPipeline(net._native(), 5, 6, True, "input", B, h, w, net.precision, depth=2)
    pipe.enable_display(shd.params22())
    outs = [torch.empty((B, 4 * h, 4 * w, 3), dtype=torch.uint8).pin_memory() for _ in range(T)]
    lows = [torch.from_numpy(np.ascontiguousarray(low[:, j])).pin_memory() for j in range(T)]
    flows = [torch.from_numpy(np.ascontiguousarray(flow[:, j])).pin_memory() for j in range(T)]
    for j in range(T):
        pipe.submit_display(lows[j], None if j == 0 else flows[j - 1], outs[j])
    pipe.wait()
    for j in range(T):
        assert np.array_equal(outs[j].numpy(), refs[j]), f"frame {j}"
    pipe.close()
