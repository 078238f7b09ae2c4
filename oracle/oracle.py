"""
oracle.py -- TEST INFRASTRUCTURE ONLY (checker for the CUDA product path).

Two CPU restatements of the reference's frame-recurrent 4x EnhanceNet inference path:

* backend "c":     ctypes binding of oracle/isr_oracle.c (plain C, fp32, explicit op order).
* backend "torch": the same functions restated with torch CPU ops (F.conv2d, F.interpolate,
                   F.grid_sample) -- the third-party arithmetic the reference itself calls
                   (PyTorch ATen, torch 2.11.0; not vendored under /root/reference).  Used for
                   sizes the scalar C code would take too long on, and as the timed CPU baseline.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing here reads /root/reference at run time.

Each function cites the reference file:line it follows (relative to
/root/reference/SuperresolutionNetwork).  Parity is pinned against tests/golden/*.npz, produced
from the live reference by oracle/gen_golden.py (the reference ships no golden vectors of its
own: "parity unpinned by the reference's tests").
"""
import ctypes
import os
import subprocess

import numpy as np
import torch
import torch.nn.functional as F

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

STATE_DICT_KEYS = (
    ["preblock.0.weight", "preblock.0.bias"]
    + [f"blocks.{i}.{j}.{k}" for i in range(10) for j in (0, 2) for k in ("weight", "bias")]
    + [f"postblock.{j}.{k}" for j in (1, 4, 6, 8) for k in ("weight", "bias")]
)


def build(force=False):
    so = os.path.join(_HERE, "libisr_oracle.so")
    src = os.path.join(_HERE, "isr_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.orc_initial_image.restype = ctypes.c_int
        _LIB.orc_enhancenet_forward.restype = ctypes.c_int
    return _LIB


def _f32(a):
    if isinstance(a, torch.Tensor):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


# ---------------------------------------------------------------------------------------------
# models/videotools.py:8-25
def flatten_high(image_high, upscale_factor, backend="c"):
    x = _f32(image_high)
    B, C, H, W = x.shape
    r = upscale_factor
    if backend == "torch":
        return F.pixel_unshuffle(torch.from_numpy(x), r).numpy()
    out = np.empty((B, C * r * r, H // r, W // r), np.float32)
    lib().orc_flatten_high(_p(x), _p(out), B, C, H, W, r)
    return out


# models/videotools.py:49-87
def warp_upscale(image_high, flow_low, upscale_factor, special_mask=False, backend="c"):
    img, flow = _f32(image_high), _f32(flow_low)
    B, C, H, W = img.shape
    _, two, h, w = flow.shape
    assert two == 2 and H == h * upscale_factor and W == w * upscale_factor
    if backend == "torch":
        ti, tf = torch.from_numpy(img), torch.from_numpy(flow)
        f2 = torch.cat((tf[:, 0:1] * -2.0, tf[:, 1:2] * 2.0), dim=1)
        fh = F.interpolate(f2, scale_factor=upscale_factor, mode="bilinear").permute(0, 2, 3, 1)
        gh = torch.linspace(-1, 1, H).unsqueeze(1)
        gw = torch.linspace(-1, 1, W).unsqueeze(0)
        offs = torch.stack(torch.broadcast_tensors(gw, gh), dim=2).unsqueeze(0)
        grid = offs + fh
        if special_mask:
            ti = torch.cat([ti[:, 0:1] * 0.5 + 0.5, ti[:, 1:]], dim=1)
        o = F.grid_sample(ti, grid, mode="bilinear", padding_mode="zeros", align_corners=False)
        if special_mask:
            o = torch.cat([o[:, 0:1] * 2 - 1, o[:, 1:]], dim=1)
        return o.numpy()
    out = np.empty_like(img)
    lib().orc_warp_upscale(_p(img), _p(flow), _p(out), B, C, h, w, upscale_factor, int(bool(special_mask)))
    return out


_MODES = {"zero": 0, "unshaded": 1, "input": 2}


# utils/initial_image.py:5-54
def initial_image(current_input, channels, mode, ao_inverted, upscaling=4):
    x = _f32(current_input)
    B, Cin, h, w = x.shape
    if mode not in _MODES:
        raise ValueError("unknown input mode: " + mode)
    out = np.empty((B, channels, h * upscaling, w * upscaling), np.float32)
    rc = lib().orc_initial_image(_p(x), _p(out), B, Cin, h, w, channels, _MODES[mode], int(bool(ao_inverted)), upscaling)
    if rc != 0:
        raise ValueError("for mode='unshaded', channels is expected to be 5 or 6")
    return out


def upsample(x, s, mode, backend="c"):
    x = _f32(x)
    B, C, h, w = x.shape
    if backend == "torch":
        return F.interpolate(torch.from_numpy(x), scale_factor=s, mode=mode).numpy()
    out = np.empty((B, C, h * s, w * s), np.float32)
    (lib().orc_upsample_bilinear if mode == "bilinear" else lib().orc_upsample_nearest)(_p(x), _p(out), B, C, h, w, s)
    return out


def conv3x3(x, weight, bias, relu=False):
    x, weight, bias = _f32(x), _f32(weight), _f32(bias)
    B, Cin, H, W = x.shape
    Cout = weight.shape[0]
    out = np.empty((B, Cout, H, W), np.float32)
    lib().orc_conv3x3(_p(x), _p(weight), _p(bias), _p(out), B, Cin, Cout, H, W, int(relu))
    return out


# models/enhancenet.py:136-145 + :51-90
def enhancenet_forward(state_dict, inputs, upsample_mode="bilinear", recon_type="residual", n_mask=5,
                       backend="c"):
    """Returns (recon, raw) like EnhanceNet.forward.  state_dict: name -> array/tensor (useBN=False)."""
    x = _f32(inputs)
    B, Cin, h, w = x.shape
    params = [_f32(state_dict[k]) for k in STATE_DICT_KEYS]
    Cout = params[-1].shape[0]
    assert params[0].shape[1] == Cin
    if upsample_mode not in ("nearest", "bilinear"):
        raise ValueError("oracle supports nearest|bilinear")
    if backend == "torch":
        with torch.no_grad():
            t = [torch.from_numpy(p) for p in params]
            xin = torch.from_numpy(x)
            f = F.relu(F.conv2d(xin, t[0], t[1], padding=1))
            pi = 2
            for _ in range(10):
                u = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1))
                u = F.conv2d(u, t[pi + 2], t[pi + 3], padding=1)
                f = f + u
                pi += 4
            f = F.interpolate(f, scale_factor=2, mode=upsample_mode)
            f = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)); pi += 2
            f = F.interpolate(f, scale_factor=2, mode=upsample_mode)
            f = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)); pi += 2
            f = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)); pi += 2
            raw = F.conv2d(f, t[pi], t[pi + 1], padding=1)
            if recon_type == "residual":
                if n_mask > Cout:
                    raise ValueError("number of output channels must be at least the number of masked input channels")
                sk = F.interpolate(xin[:, 0:n_mask], size=[raw.shape[2], raw.shape[3]], mode=upsample_mode)
                recon = torch.cat([sk + raw[:, 0:n_mask], raw[:, n_mask:]], dim=1)
            else:
                recon = raw
            return recon.numpy(), raw.numpy()
    recon = np.empty((B, Cout, 4 * h, 4 * w), np.float32)
    raw = np.empty_like(recon)
    arr = (ctypes.c_void_p * len(params))(*[p.ctypes.data for p in params])
    rc = lib().orc_enhancenet_forward(_p(x), arr, _p(recon), _p(raw), B, Cin, Cout, h, w,
                                      1 if upsample_mode == "bilinear" else 0,
                                      1 if recon_type == "residual" else 0, n_mask)
    if rc != 0:
        raise ValueError("number of output channels must be at least the number of masked input channels")
    return recon, raw


# mainVideoUnshaded.py:460-465 / mainGUI.py:594-598 / mainVideo.py:416
def postprocess(pred, unshaded=True):
    x = _f32(pred)
    B, C, H, W = x.shape
    out = np.empty_like(x)
    lib().orc_postprocess(_p(x), _p(out), B, C, H, W, int(bool(unshaded)))
    return out


# utils/shading.py:148-191
def shade(gbuffer, params):
    x = _f32(gbuffer)
    B, C, H, W = x.shape
    assert C >= 5
    prm = _f32(params)
    out = np.empty((B, 3, H, W), np.float32)
    lib().orc_shade(_p(x), _p(prm), _p(out), B, C, H, W)
    return out


# ---------------------------------------------------------------------------------------------
# Frame recurrence (mainVideoUnshaded.py:417-465 / loadedmodel.py:70-121), used as the e2e checker
# and as the timed CPU baseline.
def unshaded_sequence(state_dict, low, flow, upsample_mode="bilinear", initial="input", backend="torch",
                      shade_params=None):
    """low: B x T x 5 x h x w, flow: B x T x 2 x h x w -> list of T predictions (B x 6 x 4h x 4w)
    (and shaded RGB frames when shade_params is given)."""
    low, flow = _f32(low), _f32(flow)
    B, T = low.shape[:2]
    preds, rgbs = [], []
    prev = None
    for j in range(T):
        inp = low[:, j]
        if j == 0:
            warped = initial_image(inp, 6, initial, False, 4)
        else:
            warped = warp_upscale(prev, flow[:, j - 1], 4, special_mask=True, backend=backend)
        single = np.concatenate([inp, flatten_high(warped, 4, backend=backend)], axis=1)
        recon, _ = enhancenet_forward(state_dict, single, upsample_mode, "residual", 5, backend=backend)
        preds.append(recon)
        prev = postprocess(recon, unshaded=True)
        if shade_params is not None:
            rgbs.append(shade(prev, shade_params))
    return preds, rgbs


# The same recurrence written ONLY with torch ops, so that it can run on any torch device: on "cpu" it is the reference's
# CPU arithmetic (ATen / oneDNN), on "cuda" (TF32 off) it is what the reference computes on a GPU (ATen / cuDNN fp32).
# The difference between the two is the noise floor SURVEY.md 8(c) asks to record: no implementation can be closer to
# "the reference" than the reference's two own back ends are to each other.
def unshaded_sequence_torch(state_dict, low, flow, device="cpu", upsample_mode="bilinear", initial="input",
                            teacher=None):
    """Returns the list of T predictions (numpy).  ``teacher``: optional list of previous-frame predictions to use as
    the state instead of the run's own (teacher forcing)."""
    dev = torch.device(device)
    if dev.type == "cuda":
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
    t = [torch.from_numpy(_f32(state_dict[k])).to(dev) for k in STATE_DICT_KEYS]
    tl, tf = torch.from_numpy(_f32(low)).to(dev), torch.from_numpy(_f32(flow)).to(dev)
    B, T, _, h, w = tl.shape
    H, W = 4 * h, 4 * w

    def net(x):
        f = F.relu(F.conv2d(x, t[0], t[1], padding=1))
        pi = 2
        for _ in range(10):
            u = F.conv2d(F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)), t[pi + 2], t[pi + 3], padding=1)
            f = f + u
            pi += 4
        for _ in range(2):
            f = F.interpolate(f, scale_factor=2, mode=upsample_mode)
            f = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)); pi += 2
        f = F.relu(F.conv2d(f, t[pi], t[pi + 1], padding=1)); pi += 2
        raw = F.conv2d(f, t[pi], t[pi + 1], padding=1)
        sk = F.interpolate(x[:, 0:5], size=[H, W], mode=upsample_mode)
        return torch.cat([sk + raw[:, 0:5], raw[:, 5:]], dim=1)

    def post(p):                                               # mainVideoUnshaded.py:460-465, shading.py:193-207
        n = p[:, 1:4]
        n = n / torch.clamp(torch.norm(n, p=2, dim=1, keepdim=True), min=1e-7)
        return torch.cat([torch.clamp(p[:, 0:1], -1, 1), n, torch.clamp(p[:, 4:5], 0, 1), torch.clamp(p[:, 5:6], 0, 1)], 1)

    gh = torch.linspace(-1, 1, H, device=dev).unsqueeze(1)
    gw = torch.linspace(-1, 1, W, device=dev).unsqueeze(0)
    offs = torch.stack(torch.broadcast_tensors(gw, gh), dim=2).unsqueeze(0)
    preds = []
    prev = None
    with torch.no_grad():
        for j in range(T):
            inp = tl[:, j]
            if j == 0:
                assert initial == "input"                      # initial_image.py:39-52
                up = F.interpolate(inp, scale_factor=4, mode="bilinear")
                warped = torch.cat([up, torch.ones(B, 1, H, W, device=dev)], dim=1)
            else:
                if teacher is not None:
                    prev = post(torch.from_numpy(_f32(teacher[j - 1])).to(dev))
                f2 = torch.cat((tf[:, j - 1, 0:1] * -2.0, tf[:, j - 1, 1:2] * 2.0), dim=1)
                grid = offs + F.interpolate(f2, scale_factor=4, mode="bilinear").permute(0, 2, 3, 1)
                pm = torch.cat([prev[:, 0:1] * 0.5 + 0.5, prev[:, 1:]], dim=1)
                o = F.grid_sample(pm, grid, mode="bilinear", padding_mode="zeros", align_corners=False)
                warped = torch.cat([o[:, 0:1] * 2 - 1, o[:, 1:]], dim=1)
            pred = net(torch.cat((inp, F.pixel_unshuffle(warped, 4)), dim=1))
            preds.append(pred.cpu().numpy())
            prev = post(pred)
    return preds


# mainVideo.py:383-434 in the form the inference facade runs it (loadedmodel.py:97-119): shaded (RGB) network, previous
# RGB frame warped with special_mask=False, clamp(pred, 0, 1) as the state, zero initial image
def shaded_sequence(state_dict, inputs, flow, upsample_mode="bilinear", backend="torch"):
    """inputs: B x T x c_low x h x w (RGB, mask in [0,1][, normal][, depth]), flow: B x T x 2 x h x w
    -> list of T (unclamped) predictions B x 3 x 4h x 4w."""
    inputs, flow = _f32(inputs), _f32(flow)
    B, T, _, h, w = inputs.shape
    preds = []
    prev = None
    for j in range(T):
        inp = inputs[:, j]
        if j == 0:
            warped = initial_image(inp, 3, "zero", False, 4)                      # mainVideo.py:385-388
        else:
            warped = warp_upscale(prev, flow[:, j - 1], 4, special_mask=False, backend=backend)
        single = np.concatenate([inp, flatten_high(warped, 4, backend=backend)], axis=1)
        recon, _ = enhancenet_forward(state_dict, single, upsample_mode, "residual", 3, backend=backend)
        preds.append(recon)
        prev = postprocess(recon, unshaded=False)                                 # mainVideo.py:416
    return preds


# inference/loadedmodel.py:70-121 (one frame of the facade, including its host-side OpenCV flow in-painting -- third
# party, opencv-python, pinned 4.0.1 by the reference's Requirements.txt, 4.13 installed here)
def loadedmodel_inference(state_dict, current_low, prev_high, unshaded, initial_mode, has_normal=True, has_depth=True,
                          backend="torch"):
    import cv2 as cv
    x = _f32(current_low)
    B = x.shape[0]
    flow = np.empty((B, 2) + x.shape[2:], np.float32)
    for b in range(B):                                        # the reference handles batch 1 only ([0], :77)
        hole = np.uint8(x[b, 3] == 0)
        flow[b, 0] = cv.inpaint(x[b, 8], hole, 3, cv.INPAINT_NS)
        flow[b, 1] = cv.inpaint(x[b, 9], hole, 3, cv.INPAINT_NS)
    if unshaded:
        inp = np.concatenate([x[:, 3:4] * np.float32(2) - np.float32(1), x[:, 4:8]], axis=1)
        if prev_high is None:
            warped = initial_image(inp, 6, initial_mode, False, 4)
        else:
            warped = warp_upscale(prev_high, flow, 4, special_mask=True, backend=backend)
        n_mask = 5
    else:
        if has_normal and has_depth:
            inp = np.clip(x[:, 0:8], 0, 1)
        elif has_normal:
            inp = x[:, 0:7]
        elif has_depth:
            inp = np.concatenate([x[:, 0:4], x[:, 7:8]], axis=1)
        else:
            inp = x[:, 0:4]
        if prev_high is None:
            prev_high = initial_image(inp, 3, initial_mode, False, 4)             # :110
        warped = warp_upscale(prev_high, flow, 4, special_mask=False, backend=backend)
        n_mask = 3
    single = np.concatenate([_f32(inp), flatten_high(warped, 4, backend=backend)], axis=1)
    return enhancenet_forward(state_dict, single, "bilinear", "residual", n_mask, backend=backend)[0]


# mainGUI.py:834-849: temporal post-smoothing of the displayed RGB frame -- the previous displayed frame warped by the
# (in-painted) flow with special_mask=False, exponentially blended with the current one.  `factor` is a Python float in
# the reference; torch multiplies an fp32 tensor by fp32(factor) and by fp32(1 - factor) (1 - factor taken in double).
def temporal_blend(prev_rgb, flow, cur_rgb, factor):
    warped = warp_upscale(prev_rgb, flow, 4, special_mask=False, backend="c")
    a, b = np.float32(factor), np.float32(1.0 - float(factor))
    return (a * warped).astype(np.float32) + (b * _f32(cur_rgb)).astype(np.float32)


# Synthetic inputs come from the package's numpy-only generator so that bench / tests / oracle agree bit for bit.  The
# FILE is loaded by path: importing the package would dlopen libisr.so, and the oracle must not touch the product.
def _load_synthetic():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "isr_synthetic", os.path.join(os.path.dirname(_HERE), "isosurfacesuperresolution_b200", "synthetic.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_syn = _load_synthetic()
shading_params, synthetic_gbuffer = _syn.shading_params, _syn.synthetic_gbuffer


# mainGUI.py:603-608,615-636 (GUI.performSuperresolution, composition part; the method itself needs tkinter and cannot be
# imported, so it is restated op for op with the torch calls it makes)
def compose_frame(image_low, image_raw, shade_params=None, masking=False, background0=0.0):
    image = torch.from_numpy(_f32(image_low))
    raw = torch.from_numpy(_f32(image_raw))
    image = F.interpolate(image, scale_factor=4, mode="bilinear")                      # mainGUI.py:603 / :627
    base_mask = image[:, 3:4, :, :].clone()
    if raw.shape[1] >= 5:                                                              # unshaded, :606-608
        image[:, 0:3, :, :] = torch.from_numpy(shade(raw.numpy(), shade_params))
        image[:, 3:8, :, :] = raw[:, 0:-1, :, :]
        image[:, 10, :, :] = raw[:, -1, :, :]
    else:                                                                              # shaded, :629
        image[:, 0:3, :, :] = torch.clamp(raw, 0, 1)
    if masking:                                                                        # :630-636
        mask = base_mask * 0.5 + 0.5
        image = background0 + mask * (image - background0)
    return image.numpy()


def psnr(a, b, epsilon=1e-7):
    """utils/psnr.py:10-15 (unmasked branch): 10*log10(1/(eps + mean((a-b)^2))), peak = 1."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(10.0 * np.log10(1.0 / (epsilon + np.mean((a - b) ** 2))))


# ---- in front of the network: G-buffer ingest and flow in-painting ------------------------
INGEST_CHANNELS = {0: 5, 1: 4, 2: 5, 3: 7, 4: 8}


def ingest_gbuffer(render, variant=0, interleaved=True):
    """Channel slicing of inference/loadedmodel.py:84-105 (+ mainGUI.py:713-716 for image12) applied to the renderer's
    frame (GPURendererDirect/render_kernel.cu:254-265: RGB, mask, normal, depth, flow, AO, shadow).
    render: B x h x w x C interleaved or B x C x h x w.  Returns (low, flow, hole, image12)."""
    r = _f32(render)
    x = np.ascontiguousarray(r.transpose(0, 3, 1, 2)) if interleaved else r
    if variant == 0:
        low = np.concatenate([x[:, 3:4] * np.float32(2) - np.float32(1), x[:, 4:8]], axis=1)
    elif variant == 4:
        low = np.clip(x[:, 0:8], 0, 1)
    elif variant == 3:
        low = x[:, 0:7]
    elif variant == 2:
        low = np.concatenate([x[:, 0:4], x[:, 7:8]], axis=1)
    else:
        low = x[:, 0:4]
    hole = (x[:, 3] == 0).astype(np.uint8)
    image12 = None
    if x.shape[1] >= 12:
        image12 = x[:, :12].copy()
        image12[:, 3] = x[:, 3] * np.float32(2) - np.float32(1)
    return _f32(low), _f32(x[:, 8:10]), hole, image12


def _half_src(n_out, n_in):
    d = np.arange(n_out, dtype=np.float32)
    src = np.float32(0.5) * (d + np.float32(0.5)) - np.float32(0.5)
    src = np.maximum(src, np.float32(0))
    a = np.minimum(src.astype(np.int32), n_in - 1)
    i1 = np.where(a < n_in - 1, a + 1, a)
    l1 = np.clip(src - a.astype(np.float32), np.float32(0), np.float32(1)).astype(np.float32)
    return a, i1, (np.float32(1) - l1).astype(np.float32), l1


def inpaint_flow_pushpull(flow, hole):
    """CPU restatement of isr_inpaint_flow (isosurfacesuperresolution_b200/csrc/isr_ingest.cu), the device replacement of
    cv.inpaint(flow, mask == 0, 3, cv.INPAINT_NS) at inference/loadedmodel.py:77-82.  Not OpenCV's algorithm (push-pull
    pyramid instead of fast marching): parity with OpenCV is unpinned; this function pins the CUDA kernel's arithmetic,
    operation for operation, in fp32.  flow: B x 2 x h x w, hole: B x h x w (non-zero = fill)."""
    f = _f32(flow)
    B, _, h, w = f.shape
    wgt = (np.asarray(hole) == 0).astype(np.float32)
    vals = [np.where(wgt[:, None] > 0, f, np.float32(0)).astype(np.float32)]
    wts = [wgt]
    while vals[-1].shape[2] > 1 or vals[-1].shape[3] > 1:     # pull
        sv, sw = vals[-1], wts[-1]
        sh, swd = sv.shape[2:]
        dh, dwd = (sh + 1) // 2, (swd + 1) // 2
        W = np.zeros((B, dh, dwd), np.float32)
        V = np.zeros((B, 2, dh, dwd), np.float32)
        for dy in (0, 1):
            for dx in (0, 1):
                cw = np.zeros((B, dh, dwd), np.float32)
                cv_ = np.zeros((B, 2, dh, dwd), np.float32)
                ys, xs = sw[:, dy::2, dx::2].shape[1:]
                cw[:, :ys, :xs] = sw[:, dy::2, dx::2]
                cv_[:, :, :ys, :xs] = sv[:, :, dy::2, dx::2]
                W = (W + cw).astype(np.float32)
                V = (V + (cw[:, None] * cv_).astype(np.float32)).astype(np.float32)
        any_ = W > 0
        with np.errstate(divide="ignore", invalid="ignore"):
            dv = np.where(any_[:, None], (V / W[:, None]).astype(np.float32), np.float32(0)).astype(np.float32)
        vals.append(dv)
        wts.append(any_.astype(np.float32))
    for l in range(len(vals) - 2, -1, -1):                    # push
        cvv, fv, fw = vals[l + 1], vals[l], wts[l]
        fh, fwd = fv.shape[2:]
        y0, y1, ly0, ly1 = _half_src(fh, cvv.shape[2])
        x0, x1, lx0, lx1 = _half_src(fwd, cvv.shape[3])
        top, bot = cvv[:, :, y0], cvv[:, :, y1]
        r0 = ((lx0 * top[:, :, :, x0]).astype(np.float32) + (lx1 * top[:, :, :, x1]).astype(np.float32)).astype(np.float32)
        r1 = ((lx0 * bot[:, :, :, x0]).astype(np.float32) + (lx1 * bot[:, :, :, x1]).astype(np.float32)).astype(np.float32)
        fill = ((ly0[:, None] * r0).astype(np.float32) + (ly1[:, None] * r1).astype(np.float32)).astype(np.float32)
        vals[l] = np.where(fw[:, None] != 0, fv, fill).astype(np.float32)
        wts[l] = np.ones_like(fw)
    return vals[0]


# ---- screen-space ambient occlusion ---------------------------------------------------------
_SSAO = None
_SSAO_REF = None
_SSAO_ARGS = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
              ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
              ctypes.c_float, ctypes.c_float, ctypes.c_int]


def _ssao_lib():
    global _SSAO
    if _SSAO is None:
        so = os.path.join(_HERE, "libisr_oracle_ssao.so")
        src = os.path.join(_HERE, "ssao_oracle.cpp")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", _HERE, "-s", "libisr_oracle_ssao.so"])
        _SSAO = ctypes.CDLL(so)
        _SSAO.orc_ssao.argtypes = _SSAO_ARGS
        _SSAO.orc_ssao.restype = None
    return _SSAO


def ssao_ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libshadowfx_ref.so"))


def ssao_tables(samples=64, noise_samples=4):
    k = np.zeros(3 * samples, np.float32)
    n = np.zeros(3 * noise_samples * noise_samples, np.float32)
    _ssao_lib().orc_ssao_tables(samples, noise_samples, _p(k), _p(n))
    return k, n


def ssao(frame, samples=64, sample_radius=0.1, bias=0.025, noise_samples=4, normal_index=4, depth_index=7,
         output_index=10, use_reference_build=False):
    """ShadowFX::screenSpaceAmbientOcclusion (GPURenderer/ShadowFX.cpp:23-134) on one interleaved H x W x C frame, in the
    channel layout of its call site (GPURenderer.cpp:716-723: normal 4..6, depth 7, AO -> 10).  Returns a copy of the frame
    with the AO channel written.  ``use_reference_build``: run the reference's own ShadowFX.cpp compiled in place
    (oracle/_ref, only where /root/reference was present at build time) instead of the restatement."""
    global _SSAO_REF
    f = _f32(frame).copy()
    H, W, C = f.shape
    args = (W, H, _p(f), C, W, normal_index, normal_index + 1, normal_index + 2, depth_index, _p(f), C, W, output_index,
            samples, sample_radius, bias, noise_samples)
    if use_reference_build:
        if _SSAO_REF is None:
            _SSAO_REF = ctypes.CDLL(os.path.join(_HERE, "_ref", "libshadowfx_ref.so"))
            _SSAO_REF.ref_ssao.argtypes = _SSAO_ARGS
            _SSAO_REF.ref_ssao.restype = None
        _SSAO_REF.ref_ssao(*args)
    else:
        _ssao_lib().orc_ssao(*args)
    return f


# ---- dense-grid isosurface ray-caster ---------------------------------------------------------
_RAYCAST = None


def _raycast_lib():
    global _RAYCAST
    if _RAYCAST is None:
        so = os.path.join(_HERE, "libisr_oracle_raycast.so")
        src = os.path.join(_HERE, "raycast_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", _HERE, "-s", "libisr_oracle_raycast.so"])
        _RAYCAST = ctypes.CDLL(so)
        _RAYCAST.isr_oracle_raycast_iso.restype = None
        _RAYCAST.isr_oracle_raycast_iso.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                    ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                    ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                    ctypes.c_int]
    return _RAYCAST


def raycast_ao_tables():
    """Hemisphere samples (512 x 4) and rotation vectors (16 x 4) as GPURendererDirect.cpp:146-190 draws them."""
    hemi = np.zeros((512, 4), np.float32)
    rot = np.zeros((16, 4), np.float32)
    _ssao_lib().orc_raycast_ao_tables(_p(hemi), _p(rot))
    return hemi, rot


def raycast_iso(volume, params, width, height, ao_samples=0, hemisphere=None, rotations=None, binary_steps=10,
                viewport=None):
    """custom_iso_kernel (GPURendererDirect/render_kernel.cu:212-266) on one dense brick ``volume`` (nz x ny x nx).
    ``params``: the ISR_RC_* block of include/isr.h (81 floats).  Returns the height x width x 12 frame."""
    vol = _f32(volume)
    nz, ny, nx = vol.shape
    P = _f32(params)
    assert P.shape == (81,)
    hemi = _f32(hemisphere) if hemisphere is not None else np.zeros((max(ao_samples, 1), 4), np.float32)
    rot = _f32(rotations) if rotations is not None else np.zeros((16, 4), np.float32)
    vp = np.ascontiguousarray(viewport, dtype=np.int32) if viewport is not None else None
    out = np.empty((height, width, 12), np.float32)
    _raycast_lib().isr_oracle_raycast_iso(_p(vol), nx, ny, nz, _p(P), _p(hemi), _p(rot), int(ao_samples),
                                          int(binary_steps), _p(vp) if vp is not None else None, _p(out), width, height)
    return out
