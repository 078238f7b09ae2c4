"""Round-2 GPU parity: the shaded recurrence, the 10-frame unshaded recurrence, more weight seeds, the reference's own
LoadedModel facade (checkpoint file included), temporal post-smoothing, per-handle tunables, BatchNorm folding, the
16-bit host pipeline.  Everything goes through the C ABI; the checker is the CPU oracle / the live-reference fixtures of
oracle/gen_golden_r02.py.  Tolerances are the north star's: 1e-2 max-abs per channel and >= 50 dB PSNR."""
import os
from argparse import Namespace

import numpy as np
import pytest
import torch

from oracle import oracle as orc
from test_oracle_golden import _seeded_reference_weights, _weight_sum

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CKPT = os.path.join(ROOT, "tests", "golden", "checkpoint_unshaded_ref.pth")


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


def _net(isr, weights, cin=101, cout=6, mask=(0, 1, 2, 3, 4), ups="bilinear", bn=False):
    net = isr.createNetwork("EnhanceNet", 4, cin, list(mask), cout, Namespace(upsample=ups, reconType="residual", useBN=bn))
    if weights is not None:
        net.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in weights.items()})
    return net.cuda().eval()


def _check_net(out, ref, tol=1e-2, min_psnr=50.0):
    out = out.detach().cpu().numpy() if isinstance(out, torch.Tensor) else out
    for c in range(ref.shape[1]):
        assert float(np.max(np.abs(out[:, c] - ref[:, c]))) <= tol, f"channel {c}"
    assert orc.psnr(out, ref) >= min_psnr


# ---- the shaded (RGB) recurrence: mainVideo.py:383-434 / loadedmodel.py:97-119 -------------------------------------
@pytest.mark.parametrize("prec", ["hybrid", "bf16x3"])
def test_sequence_shaded_golden(isr, golden, prec):
    """56 -> 3 network, 3-channel previous frame (pack_recurrent_kernel<3>), special_mask=False, zero initial image, and
    clamp(pred, 0, 1) as the recurrent state written by the output conv's epilogue.  Teacher-forced against the live
    reference's frames (every frame at the bar), and the fused state against clamp of our own prediction (bit-exact)."""
    g = golden("sequence_shaded")
    sd = _seeded_reference_weights(56, 3, 1234)
    assert np.allclose(_weight_sum(sd), g["weight_sum"], rtol=1e-9)
    net = _net(isr, sd, cin=56, cout=3, mask=(0, 1, 2))
    net.precision = prec
    inputs, flow = cu(g["input"]), cu(g["flow"])
    T = inputs.shape[1]
    with torch.no_grad():
        for j in range(T):
            inp = inputs[:, j].contiguous()
            if j == 0:
                prev = isr.initialImage(inp, 3, "zero", False, 4)
                pred, _, state = net(inp, prev, None, return_state=True)
            else:
                ref_state = cu(np.clip(g[f"pred_{j - 1}"], 0, 1))
                pred, _, state = net(inp, ref_state, flow[:, j - 1].contiguous(), special_mask=False, return_state=True)
            _check_net(pred, g[f"pred_{j}"])
            assert torch.equal(state, torch.clamp(pred, 0, 1))
    # the same frames through SequenceRunner (free-running, short) stay near the reference
    runner = isr.SequenceRunner(net, initial_image="zero")
    preds, _, _ = runner.run(inputs, flow)
    _check_net(preds[0], g["pred_0"])
    assert maxabs(preds[1], g["pred_1"]) <= 3e-2


@pytest.mark.parametrize("prec", ["hybrid", "bf16x3"])
def test_sequence_unshaded_10_frames_teacher_forced(isr, golden, weights, prec):
    """configs[1] sequences have 10 frames: every frame, computed from the reference's previous output, is at the bar."""
    g = golden("sequence_unshaded_10")
    net = _net(isr, weights)
    net.precision = prec
    low, flow = cu(g["low"]), cu(g["flow"])
    worst = 0.0
    with torch.no_grad():
        for j in range(10):
            if j == 0:
                prev = isr.initialImage(low[:, 0].contiguous(), 6, "input", False, 4)
                pred, _, _ = net(low[:, 0].contiguous(), prev, None, return_state=True)
            else:
                ref_state = cu(orc.postprocess(g[f"pred_{j - 1}"], unshaded=True))
                pred, _, _ = net(low[:, j].contiguous(), ref_state, flow[:, j - 1].contiguous(), special_mask=True,
                                 return_state=True)
            _check_net(pred, g[f"pred_{j}"])
            worst = max(worst, maxabs(pred, g[f"pred_{j}"]))
    assert worst <= (1e-3 if prec == "bf16x3" else 1e-2)


def _amplification(weights, low, flow, ref, eps=1e-4, seed=0):
    """How much the network's own feedback loop amplifies a per-frame perturbation: the fp32 CPU oracle is run once more
    with uniform noise of amplitude eps added to EVERY frame's input (a stand-in for any per-frame rounding error of
    that size); gain[j] = deviation at frame j / eps.  A property of the (random-init) network, not of an implementation."""
    rng = np.random.RandomState(seed)
    noisy = low + rng.uniform(-eps, eps, low.shape).astype(np.float32)
    pert = orc.unshaded_sequence_torch(weights, noisy, flow, "cpu")
    return [float(np.max(np.abs(pert[j] - ref[j]))) / eps for j in range(len(ref))]


def test_sequence_unshaded_10_frames_free_running(isr, golden, weights):
    """Free-running over 10 frames the random-init network EXPANDS any per-frame difference.  The noise floor is measured
    right here: the reference's own arithmetic on CUDA (stock torch / cuDNN fp32, TF32 off) against the same arithmetic
    on the CPU drifts from 1.4e-5 at frame 0 to 3e-4 at frame 9 on this 16x24 fixture, and from 2.5e-5 to 1.1e-2 -- past
    the 1e-2 bar -- on 256x256 crops (profiles/r02_recurrence_drift.json).  A fixed bound on a free-running frame is
    therefore meaningless; what an implementation can be held to is that its deviation grows no faster than the
    reference's own:   err[j] <= k * growth[j] * err[0],   growth[j] = max(torch-CUDA drift[j] / drift[0], amplification
    of a per-frame input perturbation by the fp32 CPU oracle[j] / [0]),  k = 3,  err[0] = the scheme's single-frame error
    (frame 0 has no feedback), itself held to the 1e-2 bar."""
    g = golden("sequence_unshaded_10")
    ref = [g[f"pred_{j}"] for j in range(10)]
    gain = _amplification(weights, g["low"], g["flow"], ref)
    cuda_ref = orc.unshaded_sequence_torch(weights, g["low"], g["flow"], "cuda")
    floor = [maxabs(cuda_ref[j], ref[j]) for j in range(10)]
    assert 0 < floor[0] <= 1e-4                        # the two back ends of the reference differ already at frame 0
    growth = [max(floor[j] / floor[0], gain[j] / gain[0], 1.0) for j in range(10)]
    net = _net(isr, weights)
    K = 3.0
    for prec in ("hybrid", "bf16x3"):
        net.precision = prec
        preds, _, _ = isr.SequenceRunner(net, initial_image="input").run(cu(g["low"]), cu(g["flow"]))
        err = [maxabs(preds[j], ref[j]) for j in range(10)]
        assert err[0] <= 1e-2
        for j in range(10):
            assert err[j] <= K * growth[j] * err[0], (prec, j, err, growth)
        if prec == "bf16x3":                           # on this fixture the split scheme stays inside the fixed bar
            assert max(err) <= 1e-2, err


@pytest.mark.parametrize("seed", [1, 7, 42, 2024, 31337])
def test_enhancenet_more_weight_seeds(isr, golden, seed):
    """The parity margin of the default scheme is seed-dependent: five more reference-constructed networks,
    uniform stress input and a G-buffer-like input, against the live reference's outputs."""
    g = golden("enhancenet_seeds")
    sd = _seeded_reference_weights(101, 6, seed)
    assert np.allclose(_weight_sum(sd), g[f"weight_sum_{seed}"], rtol=1e-9)
    net = _net(isr, sd)
    for prec, tol in (("hybrid", 1e-2), ("bf16x3", 1e-3)):
        net.precision = prec
        for kind in ("uniform", "gbuffer"):
            with torch.no_grad():
                recon, _ = net(cu(g[f"x_{kind}"]))
            _check_net(recon, g[f"recon_{kind}_{seed}"], tol=tol)


# ---- the facade, from the reference's own checkpoint file ----------------------------------------------------------
def test_loaded_model_from_reference_checkpoint(isr, golden, weights):
    """LoadedModel(<path of a checkpoint pickled by the reference>) -> drop-in network; inference() on the reference's
    12-channel render, with the reference's host-side OpenCV in-painting, against the reference's OWN
    LoadedModel.inference outputs (first frame: initial image; second frame: warp of the post-processed state)."""
    g = golden("loadedmodel")
    lm = isr.LoadedModel(CKPT, torch.device("cuda"), 4)
    assert isinstance(lm.model, isr.EnhanceNet) and lm.unshaded and lm.initial_image_mode == "input"
    assert all(np.array_equal(v.cpu().numpy(), weights[k]) for k, v in lm.model.state_dict().items())
    p0 = lm.inference(cu(g["render_0"]), None, inpaint="opencv")
    _check_net(p0, g["unshaded_p0"])
    p1 = lm.inference(cu(g["render_1"]), cu(g["unshaded_state0"]), inpaint="opencv")
    _check_net(p1, g["unshaded_p1"])
    # the device in-painter instead of OpenCV: different fill inside the holes (parity unpinned), same frame on average
    q1 = lm.inference(cu(g["render_1"]), cu(g["unshaded_state0"]), inpaint="gpu")
    assert float((q1 - p1).abs().mean()) <= 0.05


def test_loaded_model_shaded_branch(isr, golden):
    """The shaded branch of the facade (loadedmodel.py:97-115): 8-channel clamped input, 3-channel previous frame warped
    with special_mask=False, 'zero' initial image -- against the reference's own LoadedModel.inference."""
    g = golden("loadedmodel")
    sd = _seeded_reference_weights(56, 3, 1234)
    net = _net(isr, sd, cin=56, cout=3, mask=(0, 1, 2))
    lm = isr.LoadedModel({"model": net, "parameters": {"initialImage": "zero"}}, torch.device("cuda"), 4)
    assert not lm.unshaded and lm.has_normal and lm.has_depth and lm.initial_image_mode == "zero"
    p0 = lm.inference(cu(g["render_0"]), None, inpaint="opencv")
    _check_net(p0, g["shaded_p0"])
    p1 = lm.inference(cu(g["render_1"]), cu(g["shaded_state0"]), inpaint="opencv")
    _check_net(p1, g["shaded_p1"])
    assert p1.shape == (1, 3, 64, 96)


# ---- temporal post-smoothing (mainGUI.py:834-849) ---------------------------------------------------------------------
@pytest.mark.parametrize("factor", [0.25, 0.9])
def test_temporal_smoothing_vs_oracle(isr, factor):
    rng = np.random.RandomState(5)
    _, flow, _ = orc.synthetic_gbuffer(2, 27, 48, 17)
    prev = rng.uniform(0, 1, (2, 3, 108, 192)).astype(np.float32)
    cur = rng.uniform(0, 1, (2, 3, 108, 192)).astype(np.float32)
    ref = orc.temporal_blend(prev, flow[:, 0], cur, factor)
    out = isr.temporal_smoothing(cu(prev), cu(flow[:, 0]), cu(cur), factor)
    assert maxabs(out, ref) <= 1e-5                     # the warp's tolerance; the blend itself is exact fp32
    # against the composition of the two API calls the GUI makes
    warped = isr.VideoTools.warp_upscale(cu(prev), cu(flow[:, 0]), 4)
    assert torch.equal(out, factor * warped + (1 - factor) * cu(cur))
    assert isr.temporal_smoothing(None, cu(flow[:, 0]), cu(cur), factor) is not None
    x = cu(cur)
    assert isr.temporal_smoothing(cu(prev), cu(flow[:, 0]), x, 0) is x


# ---- boundary hardening ---------------------------------------------------------------------------------------------------
def test_tunables_are_per_handle(isr, golden, weights):
    """isr_net_set_tunable changes ONE handle; isr_set_tunable only the defaults that handles created later snapshot."""
    g = golden("enhancenet_forward")
    x = cu(g["x"])
    a, b = _net(isr, weights), _net(isr, weights)
    with torch.no_grad():
        ra0, rb0 = a(x)[0].clone(), b(x)[0].clone()
        assert torch.equal(ra0, rb0)
        old = a._native().set_tunable("final_w_parts", 1)
        assert old == 0
        ra1, rb1 = a(x)[0].clone(), b(x)[0].clone()
        assert not torch.equal(ra1, ra0) and torch.equal(rb1, rb0)       # only handle `a` changed
        _check_net(ra1, g["recon_bilinear"])
        a._native().set_tunable("final_w_parts", 0)
        assert torch.equal(a(x)[0], ra0)
        # process defaults: existing handles keep their snapshot, new handles see the change
        prev = isr._lib.set_tunable("trunk_full_blocks", 4)
        try:
            assert torch.equal(b(x)[0], rb0)
            c = _net(isr, weights)
            rc = c(x)[0].clone()
            assert not torch.equal(rc, rb0)
            _check_net(rc, g["recon_bilinear"])
        finally:
            isr._lib.set_tunable("trunk_full_blocks", prev)
    with pytest.raises(isr.ops.IsrError):
        a._native().set_tunable("no_such_key", 1)


def test_two_devices_in_one_process(isr, golden, weights):
    """Kernel attributes (227 KB dynamic shared memory opt-in) are set per device: a second GPU driven from the same
    process must work.  Skipped on a one-GPU box."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    g = golden("enhancenet_forward")
    outs = []
    for d in (0, 1):
        net = _net(isr, weights).to(f"cuda:{d}")
        with torch.no_grad():
            outs.append(net(cu(g["x"]).to(f"cuda:{d}"))[0].cpu())
    assert torch.equal(outs[0], outs[1])
    _check_net(outs[1], g["recon_bilinear"])


def test_batchnorm_checkpoint_folds_into_convs(isr, weights):
    """useBN=True (enhancenet.py:99-105): eval-mode BatchNorm folds into the preceding conv; against the torch fp32
    evaluation of the same module (its differentiable path = the reference's op sequence)."""
    torch.manual_seed(1234)
    net = _net(isr, None, bn=True)
    g = torch.Generator().manual_seed(9)
    for blk in net.blocks:
        for bn in (blk[1], blk[4]):
            bn.weight.data = (torch.rand(64, generator=g) * 0.4 + 0.8).cuda()
            bn.bias.data = (torch.randn(64, generator=g) * 0.05).cuda()
            bn.running_mean = (torch.randn(64, generator=g) * 0.05).cuda()
            bn.running_var = (torch.rand(64, generator=g) * 0.4 + 0.8).cuda()
    net.invalidate()
    x = (torch.rand(1, 101, 20, 24, generator=g) * 2 - 1).cuda()
    with torch.no_grad():
        recon, raw = net(x)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    with torch.enable_grad():
        ref, _ = net._forward_autograd(x.double().float().requires_grad_(True), None, None, None)
    _check_net(recon, ref.detach().cpu().numpy())
    net.train()
    with torch.no_grad(), pytest.raises(RuntimeError):
        net(x)


def test_invalidate_after_data_update(isr, golden, weights):
    """Updates through `.data` bypass the version counter: invalidate() forces the re-pack."""
    g = golden("enhancenet_forward")
    net = _net(isr, weights)
    x = cu(g["x"])
    with torch.no_grad():
        a = net(x)[0].clone()
        net.postblock[8].bias.data.add_(0.5)
        net.invalidate()
        b = net(x)[0].clone()
    assert abs(float((b - a).mean()) - 0.5) <= 1e-3


def test_autograd_path_returns_state(isr, weights):
    net = _net(isr, weights).train()
    low, flow, _ = orc.synthetic_gbuffer(1, 12, 16, 3, T=2)
    prev = torch.rand(1, 6, 48, 64, device="cuda")
    out = net(cu(low[:, 1]).requires_grad_(True), prev, cu(flow[:, 0]), return_state=True)
    assert len(out) == 3 and out[2].shape == out[0].shape and out[0].requires_grad
    n = out[2][:, 1:4].norm(dim=1)
    assert float((n - 1).abs().max()) <= 1e-4


def test_shading_accepts_cpu_tensors_via_the_gpu(isr, golden):
    """Reference scripts shade `.cpu()` copies with ScreenSpaceShading('cpu'): the drop-in shades them with the CUDA kernel
    and returns the result on the caller's device (no CPU arithmetic)."""
    from test_gpu_parity import _shading_from_params
    g = golden("shading")
    s = _shading_from_params(isr, g["train_params"])
    out = s(torch.from_numpy(g["gbuffer"]))
    assert out.device.type == "cpu" and maxabs(out, g["train_c6"]) <= 1e-5


def test_pipeline_f16_output(isr, weights):
    """isr_pipe_submit_f16: the host receives the full prediction rounded to fp16 (half the D2H bytes); it equals the fp32
    pipeline's prediction rounded on the host, bit for bit, and the device-side recurrence is unchanged."""
    from isosurfacesuperresolution_b200 import ops
    net = _net(isr, weights)
    B, T, h, w = 2, 4, 24, 40
    low, flow, _ = orc.synthetic_gbuffer(B, h, w, 29, T=T)
    preds, _, _ = isr.SequenceRunner(net, initial_image="input").run(cu(low), cu(flow))
    pipe = ops.RecurrentPipeline(net._native(), 5, 6, True, "input", B, h, w, net.precision, depth=2)
    pipe.enable_f16()
    outs = [torch.empty((B, 6, 4 * h, 4 * w), dtype=torch.float16).pin_memory() for _ in range(T)]
    lows = [torch.from_numpy(np.ascontiguousarray(low[:, j])).pin_memory() for j in range(T)]
    flows = [torch.from_numpy(np.ascontiguousarray(flow[:, j])).pin_memory() for j in range(T)]
    for j in range(T):
        pipe.submit_f16(lows[j], None if j == 0 else flows[j - 1], outs[j])
    pipe.wait()
    for j in range(T):
        assert torch.equal(outs[j], preds[j].cpu().half()), f"frame {j}"
    pipe.close()


def test_inpaint_flow_bound_against_opencv(isr):
    """isr_inpaint_flow (push-pull) is NOT OpenCV's Navier-Stokes in-painting (loadedmodel.py:79-81) -- parity unpinned.
    This puts a number on 'visually equivalent' on the GPU box: smooth flow fields of the magnitude the renderer produces
    (|flow| <= 0.02) with silhouette-shaped holes (everything outside 3-5 spheres), error INSIDE the holes relative to the
    field's range.  Recorded in profiles/r02_inpaint_vs_opencv.json by tools/inpaint_vs_opencv.py."""
    import cv2 as cv
    from isosurfacesuperresolution_b200 import ops
    worst_mean, worst_max = 0.0, 0.0
    for seed in (1, 2, 3):
        low, flow, _ = orc.synthetic_gbuffer(1, 135, 240, seed)
        hole = (low[:, 0, 0] <= 0).astype(np.uint8)                    # outside the surface
        # the renderer's flow is smooth over the whole frame; keep only its low-frequency part (no per-pixel noise)
        yy, xx = np.meshgrid(np.linspace(-1, 1, 135, dtype=np.float32), np.linspace(-1, 1, 240, dtype=np.float32), indexing="ij")
        f = np.stack([0.012 * np.sin(2.1 * xx + seed) * np.cos(1.3 * yy), 0.012 * np.cos(1.7 * xx) * np.sin(2.3 * yy + seed)])[None]
        f = f.astype(np.float32)
        fin = f * (1 - hole[:, None])
        out = ops.inpaint_flow(cu(fin), torch.from_numpy(hole).cuda()).cpu().numpy()
        ref = np.stack([cv.inpaint(fin[0, c], hole[0], 3, cv.INPAINT_NS) for c in (0, 1)])[None]
        inside = np.broadcast_to(hole[:, None] != 0, out.shape)
        rng_f = float(f.max() - f.min())
        d = np.abs(out - ref)[inside] / rng_f
        worst_mean, worst_max = max(worst_mean, float(d.mean())), max(worst_max, float(d.max()))
        assert np.array_equal(out[~inside], fin[~inside])
        # both extrapolate the field into the hole; against the TRUE smooth field neither is far off
        assert float(np.abs(out - f)[inside].mean()) / rng_f <= 0.25
    assert worst_mean <= 0.15 and worst_max <= 0.8, (worst_mean, worst_max)


# ---- round-2 kernel variants: bit-identical to the first forms ------------------------------------------------------------
@pytest.mark.parametrize("shape", [(2, 24, 40), (1, 19, 27), (1, 33, 130), (1, 270, 480)])
@pytest.mark.parametrize("c_prev", [6, 3])
def test_pack_recurrent_v2_equals_v1(isr, weights, shape, c_prev):
    """pack_recurrent_v2_kernel (one thread per high-res column, flow rows shared by the 4 rows of a low-res row) writes
    the same network operand as the first kernel, bit for bit: the network output is compared (all planes feed it), for
    the unshaded (special mask) and the shaded (3-channel) form, on ragged widths, with flows that leave the image."""
    B, h, w = shape
    if c_prev == 6:
        net, c_low, sm = _net(isr, weights), 5, True
    else:
        torch.manual_seed(1234)
        net, c_low, sm = _net(isr, None, cin=56, cout=3, mask=(0, 1, 2)), 8, False
    g = torch.Generator(device="cuda").manual_seed(h * 100 + w)
    low = torch.rand(B, c_low, h, w, device="cuda", generator=g) * 2 - 1
    prev = torch.rand(B, c_prev, 4 * h, 4 * w, device="cuda", generator=g) * 2 - 1
    flow = (torch.rand(B, 2, h, w, device="cuda", generator=g) - 0.5) * 0.08
    flow[:, :, : h // 3] *= 20                                # far outside the image: zero / -1 padding taps
    outs = {}
    try:
        for variant in (1, 2, 3):         # 2 = v2 (default), 3 = v2 with 16-byte tap loads (experiment)
            isr._lib.set_tunable("pack_variant", variant)
            with torch.no_grad():
                outs[variant] = [net(low, prev, flow, special_mask=sm)[0].clone(), net(low, prev, None)[0].clone()]
        # a previous frame that is not 16-byte aligned takes the 4-byte path on its own
        isr._lib.set_tunable("pack_variant", 3)
        big = torch.empty(prev.numel() + 1, device="cuda")
        prev_u = big[1:].view_as(prev)
        prev_u.copy_(prev)
        with torch.no_grad():
            outs["unaligned"] = [net(low, prev_u, flow, special_mask=sm)[0].clone()]
    finally:
        isr._lib.set_tunable("pack_variant", 2)
    assert torch.equal(outs[1][0], outs[2][0]) and torch.equal(outs[1][1], outs[2][1])
    assert torch.equal(outs[1][0], outs[3][0]) and torch.equal(outs[1][0], outs["unaligned"][0])
    assert torch.isfinite(outs[2][0]).all()


@pytest.mark.parametrize("masking", [False, True])
@pytest.mark.parametrize("shape", [(1, 54, 96), (2, 19, 27), (1, 270, 480)])
def test_compose_v4_equals_scalar(isr, masking, shape):
    """compose_v4_kernel (4 output pixels per thread, 16-byte accesses) against the scalar kernel: bit-identical, for the
    unshaded 6- and 5-channel state and the shaded RGB form."""
    from test_gpu_parity import _shading_from_params
    B, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(h + w)
    img = torch.rand(B, 12, h, w, device="cuda", generator=g) * 2 - 1
    shd = _shading_from_params(isr, orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5, background=(0.2, 0.3, 0.4)))
    preds = [torch.rand(B, c, 4 * h, 4 * w, device="cuda", generator=g) * 2.4 - 1.2 for c in (6, 5, 3)]
    outs = {}
    try:
        for variant in (1, 2):
            isr._lib.set_tunable("compose_variant", variant)
            outs[variant] = [isr.compose_frame(img, p, shd if p.shape[1] >= 5 else None, masking=masking,
                                               background=(0.2, 0.3, 0.4)).clone() for p in preds]
    finally:
        isr._lib.set_tunable("compose_variant", 2)
    for a, b in zip(outs[1], outs[2]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("shape", [(1, 16, 20), (2, 19, 27), (1, 33, 130), (1, 270, 480)])
@pytest.mark.parametrize("variant", ["unshaded", "shaded", "nearest", "bf16x3"])
def test_output_conv_column_stacked_equals_tap_stacked(isr, weights, shape, variant):
    """conv3x3_final2_kernel (three kernel rows through shifted A windows, three kernel columns stacked along N, gather by
    warp shuffles; 16 x 7 tiles) against conv3x3_final_kernel (nine taps stacked, shared-memory gather; 16 x 8 tiles):
    the same products in a different fp32 summation order, prediction AND fused recurrent state, ragged sizes, Cout 6 / 3,
    bilinear / nearest skip, single and split activations."""
    B, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(3 * h + w)
    if variant == "shaded":
        torch.manual_seed(1234)
        mk = lambda: _net(isr, None, cin=56, cout=3, mask=(0, 1, 2))
        cin = 56
    else:
        mk = lambda: _net(isr, weights, ups="nearest" if variant == "nearest" else "bilinear")
        cin = 101
    x = torch.rand(B, cin, h, w, device="cuda", generator=g) * 2 - 1
    outs = {}
    try:
        for fin in (3, 1):
            isr._lib.set_tunable("final_stacked", fin)
            if variant == "shaded":
                torch.manual_seed(1234)
            net = mk()
            net.precision = "bf16x3" if variant == "bf16x3" else "hybrid"
            with torch.no_grad():
                h_ = net._native()
                outs[fin] = h_.forward(x, net.precision, want_raw=True, want_state=True, state_unshaded=(variant != "shaded"))
    finally:
        isr._lib.set_tunable("final_stacked", 3)
    for a, b in zip(outs[3], outs[1]):
        assert torch.isfinite(a).all()
        assert maxabs(a, b) <= 5e-5


@pytest.mark.parametrize("seed", [1234, 7, 2024])
def test_trunk_8bit_low_plane(isr, golden, weights, seed):
    """The default `hybrid` trunk stores the LOW activation plane as e4m3(lo * 2^6) and runs the A_lo x W correction pass as
    an 8-bit MMA (kind::f8f6f4 against e5m2(W * 2^-6), tunable trunk_lo8): half the cost of the fp16 pass.  The correction
    term only needs ~3 bits, so against the fp32 oracle the result must (1) be as good as the fp16-low-plane form, in
    max-abs and in rms, and (2) recover most of what separates that form from the scheme without any correction pass
    (K = 0) -- a wrong byte order or scale in the 8-bit path would leave it at the K = 0 level or worse.  (The two forms
    are NOT compared with each other directly: any trunk difference, however small, re-rolls the fp16 roundings of the
    tail, which puts a floor of ~5e-4 rms under such a comparison.)"""
    sd = weights if seed == 1234 else _seeded_reference_weights(101, 6, seed)
    rng = np.random.RandomState(seed)
    x = rng.uniform(-1, 1, (1, 101, 96, 128)).astype(np.float32)
    ref, _ = orc.enhancenet_forward(sd, x, "bilinear", "residual", 5, backend="torch")
    outs = {}
    try:
        # lo8: the default (8-bit low plane in the packed input, the residual stream and the conv operands);
        # lo8_convs_only: the residual stream and the packed input keep an fp16 low plane (stream_lo8 = input_lo8 = 0)
        for name, lo8, wide8, K in (("lo8", 1, 1, 10), ("lo8_convs_only", 1, 0, 10), ("lo16", 0, 0, 10), ("cheap", 0, 0, 0)):
            isr._lib.set_tunable("trunk_lo8", lo8)
            isr._lib.set_tunable("stream_lo8", wide8)
            isr._lib.set_tunable("input_lo8", wide8)
            isr._lib.set_tunable("trunk_full_blocks", K)
            net = _net(isr, sd)
            with torch.no_grad():
                outs[name] = net(cu(x))[0].cpu().numpy()
    finally:
        isr._lib.set_tunable("trunk_lo8", 1)
        isr._lib.set_tunable("stream_lo8", 1)
        isr._lib.set_tunable("input_lo8", 1)
        isr._lib.set_tunable("trunk_full_blocks", 10)
    e16, ec = (float(np.max(np.abs(outs[k] - ref))) for k in ("lo16", "cheap"))
    r16, rc = (float(np.sqrt(np.mean((outs[k] - ref) ** 2))) for k in ("lo16", "cheap"))
    assert rc >= 1.2 * r16, (r16, rc)
    for k in ("lo8", "lo8_convs_only"):
        e8 = float(np.max(np.abs(outs[k] - ref)))
        r8 = float(np.sqrt(np.mean((outs[k] - ref) ** 2)))
        assert e8 <= 1e-2 and e8 <= 1.15 * e16 + 2e-4, (k, e8, e16, ec)
        assert r8 <= 1.1 * r16, (k, r8, r16, rc)
        assert (rc - r8) >= 0.7 * (rc - r16), (k, r8, r16, rc)
        _check_net(torch.from_numpy(outs[k]), ref)


@pytest.mark.parametrize("shape", [(1, 16, 20), (2, 19, 27), (1, 33, 130), (1, 270, 480)])
def test_fused_upsample_wide_equals_four_phase(isr, weights, shape):
    """The fused upsample + conv layers issue the two x-phases of an output row as ONE N = 128 MMA operand (tunable
    up2_wide = 1, the default: conv3x3_pair_kernel<0, 64, 0, 2>) instead of four N = 64 phase convolutions (up2_wide = 0):
    the same products accumulated in the same order, so the network output is bit-identical -- on ragged tiles, on a
    frame with an odd number of tile pairs and at 1080p / 4."""
    from isosurfacesuperresolution_b200 import _lib
    B, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(17)
    x = torch.rand(B, 101, h, w, device="cuda", generator=g) * 2 - 1
    outs = {}
    try:
        for wide in (1, 0):
            _lib.set_tunable("up2_wide", wide)
            net = _net(isr, weights)
            with torch.no_grad():
                outs[wide] = net(x)[0].clone()
    finally:
        _lib.set_tunable("up2_wide", 1)
    assert torch.isfinite(outs[1]).all()
    assert torch.equal(outs[1], outs[0])


@pytest.mark.parametrize("shape", [(1, 16, 20), (2, 19, 27), (3, 33, 130), (1, 270, 480)])
def test_alternating_tile_order_is_bit_identical(isr, weights, shape):
    """Consecutive convs walk their tiles in alternating directions (tunable alt_order = 1, the default; ConvParams::reverse)
    so that each launch starts on the part of its input that is still in L2.  The order in which persistent CTAs visit the
    tiles cannot change a result: bit-identical to alt_order = 0 on ragged tiles, odd tile counts (the unpaired last tile
    is the FIRST one a reversed pair kernel visits) and several images."""
    from isosurfacesuperresolution_b200 import _lib
    B, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(23)
    x = torch.rand(B, 101, h, w, device="cuda", generator=g) * 2 - 1
    outs = {}
    try:
        for alt in (1, 0):
            _lib.set_tunable("alt_order", alt)
            net = _net(isr, weights)
            with torch.no_grad():
                outs[alt] = net(x)[0].clone()
    finally:
        _lib.set_tunable("alt_order", 1)
    assert torch.isfinite(outs[1]).all()
    assert torch.equal(outs[1], outs[0])
