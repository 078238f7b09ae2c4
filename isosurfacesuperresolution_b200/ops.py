"""Torch-tensor front end of the C ABI: device memory and streams come from PyTorch, the arithmetic from libisr.so.

Every function requires CUDA fp32 tensors and raises otherwise -- there is no CPU path.
"""
import ctypes
import numpy as np

import torch

from . import _lib
from ._lib import INITIAL, PREC, UPSAMPLE, IsrError, check, lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _prep(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: the B200 path needs CUDA tensors (got {t.device}); there is no CPU fallback")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def flatten_high(image_high, upscale_factor):
    """VideoTools.flatten_high (models/videotools.py:8-25)."""
    x = _prep(image_high, "image_high")
    B, C, H, W = x.shape
    r = int(upscale_factor)
    assert H % r == 0 and W % r == 0
    out = torch.empty((B, C * r * r, H // r, W // r), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(lib().isr_flatten_high(_p(x), _p(out), B, C, H, W, r, _stream()))
    return out


def warp_upscale(image_high, flow_low, upscale_factor, special_mask=False, align_corners=False):
    """VideoTools.warp_upscale (models/videotools.py:49-87)."""
    img, flow = _prep(image_high, "image_high"), _prep(flow_low, "flow_low")
    B, C, H, W = img.shape
    Bf, two, h, w = flow.shape
    r = int(upscale_factor)
    assert two == 2 and Bf == B and H == h * r and W == w * r     # videotools.py:63
    out = torch.empty_like(img)
    with torch.cuda.device(img.device):
        check(lib().isr_warp_upscale(_p(img), _p(flow), _p(out), B, C, h, w, r, int(bool(special_mask)),
                                     int(bool(align_corners)), _stream()))
    return out


def temporal_blend(prev_rgb, flow_low, cur_rgb, factor, upscale_factor=4):
    """Temporal post-smoothing of the displayed frame (mainGUI.py:834-849) in one pass:
    ``factor * warp_upscale(prev_rgb, flow, special_mask=False) + (1 - factor) * cur_rgb``."""
    prev, flow, cur = _prep(prev_rgb, "prev_rgb"), _prep(flow_low, "flow_low"), _prep(cur_rgb, "cur_rgb")
    B, C, H, W = cur.shape
    r = int(upscale_factor)
    assert tuple(prev.shape) == tuple(cur.shape) and tuple(flow.shape) == (B, 2, H // r, W // r) and H % r == 0 and W % r == 0
    out = torch.empty_like(cur)
    with torch.cuda.device(cur.device):
        check(lib().isr_temporal_blend(_p(prev), _p(flow), _p(cur), _p(out), B, C, H // r, W // r, r, float(factor),
                                       _stream()))
    return out


def initial_image(current_input, channels, mode, ao_inverted, upscaling=4):
    """initialImage (utils/initial_image.py:5-54)."""
    if mode not in INITIAL:
        raise ValueError("unknown input mode: " + str(mode))
    if mode == "unshaded" and channels not in (5, 6):
        raise ValueError("for mode='unshaded', channels is expected to be 5 or 6")
    x = _prep(current_input, "current_input")
    B, Cin, h, w = x.shape
    out = torch.empty((B, channels, h * upscaling, w * upscaling), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(lib().isr_initial_image(_p(x), _p(out), B, Cin, h, w, channels, INITIAL[mode], int(bool(ao_inverted)),
                                      int(upscaling), _stream()))
    return out


def postprocess(pred, unshaded=True):
    """clamp / normalise between frames (mainVideoUnshaded.py:460-465, mainVideo.py:416)."""
    x = _prep(pred, "pred")
    B, C, H, W = x.shape
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        check(lib().isr_postprocess(_p(x), _p(out), B, C, H, W, int(bool(unshaded)), _stream()))
    return out


def shade(gbuffer, params22):
    """ScreenSpaceShading.forward (utils/shading.py:148-191); params22: 22 python floats (include/isr.h)."""
    x = _prep(gbuffer, "input")
    B, C, H, W = x.shape
    assert C >= 5                                                   # shading.py:150
    arr = (ctypes.c_float * 22)(*[float(v) for v in params22])
    out = torch.empty((B, 3, H, W), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(lib().isr_shade(_p(x), arr, _p(out), B, C, H, W, _stream()))
    return out


def shade_rgb8(gbuffer, params22):
    """isr_shade_rgb8: shading straight to B x H x W x 3 uint8 (uint8(clip(rgb * 255, 0, 255)), mainGUI.py:507)."""
    x = _prep(gbuffer, "input")
    B, C, H, W = x.shape
    assert C >= 5
    arr = (ctypes.c_float * 22)(*[float(v) for v in params22])
    out = torch.empty((B, H, W, 3), device=x.device, dtype=torch.uint8)
    with torch.cuda.device(x.device):
        check(lib().isr_shade_rgb8(_p(x), arr, _p(out), B, C, H, W, _stream()))
    return out


def compose(image_low, pred, params22, unshaded=True, masking=False, background0=0.0):
    """GUI / video composition (mainGUI.py:603-608,615-636): 12-channel bilinear x4 + overwrite + optional mask lerp."""
    low, pr = _prep(image_low, "image_low"), _prep(pred, "pred")
    B, c_low, h, w = low.shape
    assert tuple(pr.shape[2:]) == (4 * h, 4 * w) and pr.shape[0] == B
    arr = (ctypes.c_float * 22)(*[float(v) for v in params22]) if params22 is not None else None
    out = torch.empty((B, c_low, 4 * h, 4 * w), device=low.device, dtype=torch.float32)
    with torch.cuda.device(low.device):
        check(lib().isr_compose(_p(low), _p(pr), arr, _p(out), B, c_low, pr.shape[1], h, w, int(bool(unshaded)),
                                int(bool(masking)), float(background0), _stream()))
    return out


INGEST_VARIANTS = {"unshaded": 0, "shaded": 1, "shaded_depth": 2, "shaded_normal": 3, "shaded_normal_depth": 4}


def ingest_gbuffer(render, variant="unshaded", interleaved=None, want_flow=True, want_hole=True, want_image12=False):
    """One-pass G-buffer ingest (include/isr.h isr_ingest_gbuffer).  ``render``: the renderer's frame, H x W x 12 or
    B x H x W x 12 interleaved (GPURendererDirect wire format; ``interleaved=True``) or planar B x C x h x w (the tensor
    LoadedModel.inference receives).  ``interleaved=None`` guesses: last dimension 12 and second dimension != 12.
    Returns (low, flow, hole, image12); entries not asked for are None."""
    x = _prep(render, "render")
    if x.dim() == 3:
        x = x.unsqueeze(0)
    if interleaved is None:
        interleaved = x.shape[-1] in (10, 12) and x.shape[1] not in (10, 12)
    if interleaved:
        B, h, w, c = x.shape
    else:
        B, c, h, w = x.shape
    v = INGEST_VARIANTS[variant] if isinstance(variant, str) else int(variant)
    c_low = check(lib().isr_ingest_low_channels(v))
    dev = x.device
    low = torch.empty((B, c_low, h, w), device=dev, dtype=torch.float32)
    flow = torch.empty((B, 2, h, w), device=dev, dtype=torch.float32) if want_flow else None
    hole = torch.empty((B, h, w), device=dev, dtype=torch.uint8) if want_hole else None
    img = torch.empty((B, 12, h, w), device=dev, dtype=torch.float32) if want_image12 else None
    with torch.cuda.device(dev):
        check(lib().isr_ingest_gbuffer(_p(x), int(bool(interleaved)), B, c, h, w, v, _p(low), _p(flow), _p(hole), _p(img),
                                       _stream()))
    return low, flow, hole, img


def inpaint_flow(flow, hole):
    """Device flow in-painting (push-pull pyramid), the replacement of cv.inpaint(..., 3, INPAINT_NS) in
    loadedmodel.py:77-82.  flow: B x 2 x h x w, hole: B x h x w (non-zero = to be filled)."""
    f = _prep(flow, "flow")
    if not isinstance(hole, torch.Tensor) or not hole.is_cuda:
        raise RuntimeError("hole: the B200 path needs a CUDA tensor; there is no CPU fallback")
    hm = hole.contiguous() if hole.dtype == torch.uint8 else (hole != 0).to(torch.uint8).contiguous()
    B, two, h, w = f.shape
    assert two == 2 and tuple(hm.shape) == (B, h, w)
    nbytes = _lib._sz(0)
    check(lib().isr_inpaint_workspace_bytes(B, h, w, ctypes.byref(nbytes)))
    ws = torch.empty(nbytes.value + 16, device=f.device, dtype=torch.uint8)
    off = (-ws.data_ptr()) % 16
    out = torch.empty_like(f)
    with torch.cuda.device(f.device):
        check(lib().isr_inpaint_flow(_p(f), _p(hm), _p(out), B, h, w, ctypes.c_void_p(ws.data_ptr() + off), nbytes.value,
                                     _stream()))
    return out


def ssao(frame, samples=64, sample_radius=0.1, bias=0.025, noise_samples=4, normal_index=4, depth_index=7,
         output_index=10):
    """ShadowFX::screenSpaceAmbientOcclusion (GPURenderer/ShadowFX.cpp:23-134) on the device, IN PLACE on the renderer's
    interleaved frame (H x W x C or B x H x W x C CUDA tensor; channel layout of GPURenderer.cpp:716-723).  Returns it."""
    if not isinstance(frame, torch.Tensor) or not frame.is_cuda or frame.dtype != torch.float32 or not frame.is_contiguous():
        raise RuntimeError("frame: a contiguous CUDA float32 tensor is required; there is no CPU fallback")
    views = [frame] if frame.dim() == 3 else list(frame.unbind(0))
    with torch.cuda.device(frame.device):
        for f in views:
            H, W, C = f.shape
            check(lib().isr_ssao(W, H, _p(f), C, W, normal_index, normal_index + 1, normal_index + 2, depth_index, _p(f),
                                 C, W, output_index, int(samples), float(sample_radius), float(bias), int(noise_samples),
                                 _stream()))
    return frame


RC_PARAMS = 81   # include/isr.h ISR_RC_PARAMS


def raycast_ao_tables():
    """Hemisphere samples (512 x 4) and rotation vectors (16 x 4) of the ray-caster's ambient occlusion, drawn as
    GPURendererDirect.cpp:146-190 draws them (host arrays)."""
    hemi = np.zeros((512, 4), np.float32)
    rot = np.zeros((16, 4), np.float32)
    check(lib().isr_raycast_ao_tables(hemi.ctypes.data_as(ctypes.c_void_p), rot.ctypes.data_as(ctypes.c_void_p)))
    return hemi, rot


def raycast_build_skip(volume):
    """Empty-space grid of a brick for ``raycast_iso(..., skip=...)``: one float per 8^3 macro-cell (device tensor)."""
    if not isinstance(volume, torch.Tensor) or not volume.is_cuda or volume.dtype != torch.float32 or not volume.is_contiguous():
        raise RuntimeError("volume: a contiguous CUDA float32 tensor is required; there is no CPU fallback")
    nz, ny, nx = volume.shape
    n = ctypes.c_size_t(0)
    check(lib().isr_raycast_skip_bytes(nx, ny, nz, ctypes.byref(n)))
    skip = torch.empty(n.value // 4, device=volume.device, dtype=torch.float32)
    with torch.cuda.device(volume.device):
        check(lib().isr_raycast_build_skip(_p(volume), nx, ny, nz, _p(skip), n.value, _stream()))
    return skip


def raycast_iso(volume, params, out, ao_samples=0, hemisphere=None, rotations=None, binary_steps=10, viewport=None,
                skip=None):
    """custom_iso_kernel (GPURendererDirect/render_kernel.cu:212-266) on one dense brick: ``volume`` nz x ny x nx CUDA
    float32, ``params`` the 81-float ISR_RC_* block (host), ``out`` a caller-allocated H x W x 12 CUDA float32 tensor
    (the reference's render(target) convention, inference/renderer.py:104-107).  Asynchronous on the current stream."""
    for name, t in (("volume", volume), ("out", out)):
        if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError(f"{name}: a contiguous CUDA float32 tensor is required; there is no CPU fallback")
    assert volume.dim() == 3 and out.dim() == 3 and out.shape[2] == 12
    P = np.ascontiguousarray(params, dtype=np.float32)
    assert P.shape == (RC_PARAMS,)
    nz, ny, nx = volume.shape
    H, W, _ = out.shape
    hemi = np.ascontiguousarray(hemisphere, dtype=np.float32) if hemisphere is not None else None
    rot = np.ascontiguousarray(rotations, dtype=np.float32) if rotations is not None else None
    assert hemi is None or hemi.size >= 4 * ao_samples
    assert rot is None or rot.size == 64
    vp = np.ascontiguousarray(viewport, dtype=np.int32) if viewport is not None else None
    vpp = lambda a: a.ctypes.data_as(ctypes.c_void_p) if a is not None else None
    with torch.cuda.device(out.device):
        check(lib().isr_raycast_iso(_p(volume), nx, ny, nz, vpp(P), vpp(hemi), vpp(rot), int(ao_samples), int(binary_steps),
                                    vpp(vp), _p(skip) if skip is not None else None, _p(out), W, H, _stream()))
    return out


def _workspace(device, nbytes):
    # cudaMalloc'd blocks from torch's caching allocator are 512-byte aligned; over-allocate and slice to 1024
    buf = torch.empty(nbytes + 1024, device=device, dtype=torch.uint8)
    off = (-buf.data_ptr()) % 1024
    return buf, buf.data_ptr() + off


def debug_conv3x3(x, weight, bias=None, addend=None, relu=False, fmt="bf16", n_act_parts=1, n_w_parts=1):
    """One tensor-core conv3x3 (64 -> cout) with fp32 NCHW in/out -- unit-test hook."""
    x = _prep(x, "x")
    weight = _prep(weight, "weight")
    B, C, H, W = x.shape
    cout = weight.shape[0]
    assert C == 64 and tuple(weight.shape[1:]) == (64, 3, 3)
    bias = _prep(bias, "bias") if bias is not None else None
    addend = _prep(addend, "addend") if addend is not None else None
    out = torch.empty((B, cout, H, W), device=x.device, dtype=torch.float32)
    n = ctypes.c_size_t(0)
    check(lib().isr_debug_conv_workspace_bytes(B, H, W, ctypes.byref(n)))
    buf, ptr = _workspace(x.device, n.value)
    with torch.cuda.device(x.device):
        check(lib().isr_debug_conv3x3(_p(x), _p(weight), _p(bias), _p(addend), _p(out), B, H, W, cout, int(bool(relu)),
                                      0 if fmt == "bf16" else 1, n_act_parts, n_w_parts, ctypes.c_void_p(ptr),
                                      n.value, _stream()))
        torch.cuda.current_stream().synchronize()   # buf must outlive the kernels
    return out


class NetHandle:
    """Owns an isr_net (packed weights + cached execution plans) and the activation workspace."""

    def __init__(self, input_channels, output_channels, n_mask, upsample, recon_residual, device):
        if upsample not in UPSAMPLE:
            raise ValueError("the B200 kernels support upsample = nearest | bilinear, got " + str(upsample))
        self.device = torch.device(device)
        self.cin, self.cout = int(input_channels), int(output_channels)
        self._net = ctypes.c_void_p(0)
        with torch.cuda.device(self.device):
            check(lib().isr_net_create(ctypes.byref(self._net), self.cin, self.cout, int(n_mask), UPSAMPLE[upsample],
                                       int(bool(recon_residual))))
        self._ws = {}          # (B, h, w, precision) -> (tensor, aligned pointer, bytes), least recently used first

    def set_tunable(self, key, value):
        """Changes a tunable of THIS handle (include/isr.h isr_net_set_tunable); returns the previous value."""
        return check(lib().isr_net_set_tunable(self._net, key.encode(), int(value)))

    def __del__(self):
        try:
            if self._net:
                lib().isr_net_destroy(self._net)
                self._net = ctypes.c_void_p(0)
        except Exception:
            pass

    def load_weights(self, tensors):
        ts = [_prep(t.detach(), "parameter") for t in tensors]
        assert len(ts) == 50
        arr = (ctypes.c_void_p * 50)(*[t.data_ptr() for t in ts])
        with torch.cuda.device(self.device):
            # (re)packing overwrites the buffers every plan reads: drain all streams of the device first -- a
            # RecurrentPipeline or a forward on a side stream may still be running with the old weights
            torch.cuda.synchronize(self.device)
            check(lib().isr_net_load_weights(self._net, arr, 50, _stream()))

    def workspace(self, B, h, w, precision):
        key = (B, h, w, precision)
        if key in self._ws:
            self._ws[key] = self._ws.pop(key)            # most recently used last
            return self._ws[key]
        n = ctypes.c_size_t(0)
        check(lib().isr_net_workspace_bytes(self._net, B, h, w, PREC[precision], ctypes.byref(n)))
        if len(self._ws) >= 4:
            # evict the least recently used block -- only after every stream of the device has drained: kernels of a
            # forward issued on a non-current stream may still be using it, and torch's caching allocator would hand the
            # block to the next tensor of the current stream at once.  (RecurrentPipeline keeps its own reference.)
            torch.cuda.synchronize(self.device)
            self._ws.pop(next(iter(self._ws)))
        buf, ptr = _workspace(self.device, n.value)
        self._ws[key] = (buf, ptr, n.value)
        return self._ws[key]

    def forward(self, inputs, precision="hybrid", want_raw=True, want_state=False, state_unshaded=True):
        x = _prep(inputs, "inputs")
        B, C, h, w = x.shape
        assert C == self.cin
        _, ptr, nbytes = self.workspace(B, h, w, precision)
        shape = (B, self.cout, 4 * h, 4 * w)
        recon = torch.empty(shape, device=x.device, dtype=torch.float32)
        raw = torch.empty(shape, device=x.device, dtype=torch.float32) if want_raw else None
        state = torch.empty(shape, device=x.device, dtype=torch.float32) if want_state else None
        with torch.cuda.device(x.device):
            check(lib().isr_net_forward(self._net, _p(x), _p(recon), _p(raw), _p(state), int(bool(state_unshaded)), B, h,
                                        w, PREC[precision], ctypes.c_void_p(ptr), nbytes, _stream()))
        return recon, raw, state

    def forward_recurrent(self, low, prev, flow, special_mask, precision="hybrid", align_corners=False, want_raw=False,
                          want_state=True, state_unshaded=True, out_state=None):
        low, prev = _prep(low, "low"), _prep(prev, "previous_output")
        flow = _prep(flow, "flow") if flow is not None else None
        B, c_low, h, w = low.shape
        c_prev = prev.shape[1]
        assert tuple(prev.shape) == (B, c_prev, 4 * h, 4 * w)
        _, ptr, nbytes = self.workspace(B, h, w, precision)
        shape = (B, self.cout, 4 * h, 4 * w)
        recon = torch.empty(shape, device=low.device, dtype=torch.float32)
        raw = torch.empty(shape, device=low.device, dtype=torch.float32) if want_raw else None
        state = None
        if want_state:
            state = out_state if out_state is not None else torch.empty(shape, device=low.device, dtype=torch.float32)
        with torch.cuda.device(low.device):
            check(lib().isr_net_forward_recurrent(self._net, _p(low), _p(prev), _p(flow), c_low, c_prev,
                                                  int(bool(special_mask)), int(bool(align_corners)), _p(recon), _p(raw),
                                                  _p(state), int(bool(state_unshaded)), B, h, w, PREC[precision],
                                                  ctypes.c_void_p(ptr), nbytes, _stream()))
        return recon, raw, state

    def time_conv(self, B, h, w, precision="hybrid", scale=4, iters=10):
        _, ptr, nbytes = self.workspace(B, h, w, precision)
        ms = ctypes.c_float(0)
        with torch.cuda.device(self.device):
            check(lib().isr_net_time_conv(self._net, B, h, w, PREC[precision], scale, ctypes.c_void_p(ptr), nbytes,
                                          iters, ctypes.byref(ms), _stream()))
        return float(ms.value)


class RecurrentPipeline:
    """Software-pipelined frame recurrence with HOST input / output buffers (include/isr.h ``isr_pipe_*``): the H2D copy
    of frame j+1 and the D2H copy of prediction j overlap the kernels of frame j+1; the recurrent state stays on the
    device.  Host tensors should be pinned (``tensor.pin_memory()``), otherwise the copies serialise."""

    def __init__(self, handle, c_low, c_prev, special_mask, initial_mode, B, h, w, precision="hybrid", depth=2):
        if initial_mode not in INITIAL:
            raise ValueError("unknown input mode: " + str(initial_mode))
        self.handle, self.B, self.h, self.w, self.c_low = handle, int(B), int(h), int(w), int(c_low)
        dev = handle.device
        # the pipeline launches on its own streams: it owns a reference to the workspace tensor (an eviction from the
        # handle's cache must not free it under running kernels) and allocates nothing from the caching allocator later
        self._ws_tensor, ws_ptr, ws_bytes = handle.workspace(B, h, w, precision)
        n = ctypes.c_size_t(0)
        check(lib().isr_pipe_staging_floats(handle._net, c_low, B, h, w, depth, ctypes.byref(n)))
        self.state = torch.zeros((B, c_prev, 4 * h, 4 * w), device=dev, dtype=torch.float32)
        self._staging = torch.empty(n.value, device=dev, dtype=torch.float32)
        self.out_shape = (B, handle.cout, 4 * h, 4 * w)
        self._pipe = ctypes.c_void_p(0)
        torch.cuda.synchronize(dev)      # the pipeline runs on its own streams
        with torch.cuda.device(dev):
            check(lib().isr_pipe_create(ctypes.byref(self._pipe), handle._net, c_low, c_prev, int(bool(special_mask)),
                                        INITIAL[initial_mode], B, h, w, PREC[precision], ctypes.c_void_p(ws_ptr), ws_bytes,
                                        _p(self.state), _p(self._staging), n.value, depth))

    @staticmethod
    def _host(t, name, shape=None):
        if not isinstance(t, torch.Tensor) or t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
            raise TypeError(f"{name}: expected a contiguous float32 CPU tensor (pinned for overlap)")
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"{name}: expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return ctypes.c_void_p(t.data_ptr())

    def submit(self, low_host, flow_host, recon_host):
        """Enqueue one frame; ``flow_host=None`` starts a new sequence.  ``recon_host`` is valid after ``wait()``."""
        lo = self._host(low_host, "low_host", (self.B, self.c_low, self.h, self.w))
        fl = self._host(flow_host, "flow_host", (self.B, 2, self.h, self.w)) if flow_host is not None else ctypes.c_void_p(0)
        out = self._host(recon_host, "recon_host", self.out_shape)
        with torch.cuda.device(self.handle.device):
            check(lib().isr_pipe_submit(self._pipe, lo, fl, out))

    def enable_display(self, params22):
        """Switch on the display form: ``submit_display`` returns shaded 8-bit RGB frames (B x 4h x 4w x 3) to the host."""
        arr = (ctypes.c_float * 22)(*[float(v) for v in params22])
        depth_bytes = 2 * self.B * 16 * self.h * self.w * 3 * 4      # room for depth <= 8
        self._rgb8 = torch.empty(depth_bytes, device=self.handle.device, dtype=torch.uint8)
        check(lib().isr_pipe_enable_display(self._pipe, arr, _p(self._rgb8), depth_bytes))

    def submit_display(self, low_host, flow_host, rgb8_host):
        lo = self._host(low_host, "low_host", (self.B, self.c_low, self.h, self.w))
        fl = self._host(flow_host, "flow_host", (self.B, 2, self.h, self.w)) if flow_host is not None else ctypes.c_void_p(0)
        if (not isinstance(rgb8_host, torch.Tensor) or rgb8_host.is_cuda or rgb8_host.dtype != torch.uint8
                or tuple(rgb8_host.shape) != (self.B, 4 * self.h, 4 * self.w, 3) or not rgb8_host.is_contiguous()):
            raise TypeError("rgb8_host: expected a contiguous uint8 CPU tensor of shape B x 4h x 4w x 3 (pinned for overlap)")
        with torch.cuda.device(self.handle.device):
            check(lib().isr_pipe_submit_display(self._pipe, lo, fl, ctypes.c_void_p(rgb8_host.data_ptr())))

    def set_state(self, state):
        """Use another recurrent-state tensor (B x c_prev x 4h x 4w, CUDA fp32) for the frames submitted from now on: one
        pipeline serves several chunks of independent sequences in turn (configs[4]: 64 sequences as 16 chunks of 4)."""
        if (not isinstance(state, torch.Tensor) or not state.is_cuda or state.dtype != torch.float32
                or tuple(state.shape) != tuple(self.state.shape) or not state.is_contiguous()):
            raise TypeError("state: expected a contiguous CUDA float32 tensor of shape " + str(tuple(self.state.shape)))
        if not hasattr(self, "_states"):
            self._states = [self.state]        # frames in flight on the pipeline's own streams may still use them
        if all(t is not state for t in self._states):
            self._states.append(state)
        self.state = state
        check(lib().isr_pipe_set_state(self._pipe, _p(state)))

    def enable_f16(self):
        """Switch on the 16-bit form: ``submit_f16`` returns the full prediction rounded to fp16 (half the D2H bytes)."""
        n = 1
        for d in self.out_shape:
            n *= d
        nbytes = 8 * n * 2                                            # room for depth <= 8
        self._f16 = torch.empty(nbytes, device=self.handle.device, dtype=torch.uint8)
        check(lib().isr_pipe_enable_f16(self._pipe, _p(self._f16), nbytes))

    def submit_f16(self, low_host, flow_host, recon_f16_host):
        lo = self._host(low_host, "low_host", (self.B, self.c_low, self.h, self.w))
        fl = self._host(flow_host, "flow_host", (self.B, 2, self.h, self.w)) if flow_host is not None else ctypes.c_void_p(0)
        if (not isinstance(recon_f16_host, torch.Tensor) or recon_f16_host.is_cuda or recon_f16_host.dtype != torch.float16
                or tuple(recon_f16_host.shape) != tuple(self.out_shape) or not recon_f16_host.is_contiguous()):
            raise TypeError("recon_f16_host: expected a contiguous float16 CPU tensor of the prediction's shape (pinned for overlap)")
        with torch.cuda.device(self.handle.device):
            check(lib().isr_pipe_submit_f16(self._pipe, lo, fl, ctypes.c_void_p(recon_f16_host.data_ptr())))

    def wait(self):
        check(lib().isr_pipe_wait(self._pipe))

    def close(self):
        if self._pipe:
            lib().isr_pipe_destroy(self._pipe)
            self._pipe = ctypes.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


__all__ = ["RecurrentPipeline", "compose", "temporal_blend", "flatten_high", "warp_upscale", "initial_image", "postprocess", "shade", "debug_conv3x3", "NetHandle",
           "IsrError", "PREC"]
