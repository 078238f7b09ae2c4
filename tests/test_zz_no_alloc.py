"""isr_inpaint_flow_static: the allocation-free variant of the flow in-painting entry point (include/isr.h), whose pyramid
scratch is a static device array inside libisr.so.  It has no counterpart in the reference (added
for callers that cannot call cudaMalloc); the checker is the existing entry point isr_inpaint_flow, which
runs the same kernels on a caller workspace.  Collected last on purpose.

Everything goes through the C ABI (ctypes); torch only provides the device buffers of the inputs and outputs."""
import ctypes

import numpy as np
import pytest
import torch

ISR_ERR_WORKSPACE = -3
MAX_BYTES = 8 << 20          # ISR_INPAINT_STATIC_MAX_BYTES
TOL = 1e-7                   # the tolerance of the in-painting tests (tests/test_ingest_oracle.py, test_gpu_parity.py)


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _case(B, h, w, seed):
    """A smooth flow with per-pixel noise, silhouette-like holes (discs) plus scattered single-pixel holes."""
    rng = np.random.RandomState(seed)
    yy, xx = np.meshgrid(np.linspace(-1, 1, h), np.linspace(-1, 1, w), indexing="ij")
    flow = np.stack([0.02 * np.sin(2.1 * xx + seed) * np.cos(1.3 * yy), 0.02 * np.cos(1.7 * xx) * np.sin(2.3 * yy + seed)])
    flow = np.tile(flow[None], (B, 1, 1, 1)) + rng.uniform(-0.004, 0.004, (B, 2, h, w))
    hole = rng.uniform(size=(B, h, w)) < 0.05
    for b in range(B):
        for _ in range(4):
            cy, cx, r = rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(0.1, 0.45)
            hole[b] |= (yy - cy) ** 2 + (xx - cx) ** 2 < r * r
    return flow.astype(np.float32), hole.astype(np.uint8)


def _workspace_call(lib, flow, hole):
    B, _, h, w = flow.shape
    nbytes = ctypes.c_size_t(0)
    assert lib.isr_inpaint_workspace_bytes(B, h, w, ctypes.byref(nbytes)) == 0
    ws = torch.empty(nbytes.value + 16, device="cuda", dtype=torch.uint8)
    off = (-ws.data_ptr()) % 16
    out = torch.full_like(flow, float("nan"))
    rc = lib.isr_inpaint_flow(_vp(flow), _vp(hole), _vp(out), B, h, w, ctypes.c_void_p(ws.data_ptr() + off), nbytes.value,
                              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.isr_last_error().decode()
    torch.cuda.synchronize()
    return out, nbytes.value


def _static_call(lib, flow, hole):
    B, _, h, w = flow.shape
    out = torch.full_like(flow, float("nan"))
    rc = lib.isr_inpaint_flow_static(_vp(flow), _vp(hole), _vp(out), B, h, w,
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.isr_last_error().decode()
    torch.cuda.synchronize()
    return out


def test_static_inpaint_refuses_sizes_beyond_the_built_in_scratch():
    """Sizes whose pyramid does not fit the 8 MiB array are refused with ISR_ERR_WORKSPACE and a message that names the
    limit -- before any device work, so this needs no GPU (the pointers are never dereferenced)."""
    from isosurfacesuperresolution_b200 import _lib
    lib = _lib.lib()
    dummy = (ctypes.c_float * 4)()
    p = ctypes.c_void_p(ctypes.addressof(dummy))
    need = ctypes.c_size_t(0)
    assert lib.isr_inpaint_workspace_bytes(4, 512, 513, ctypes.byref(need)) == 0 and need.value > MAX_BYTES
    rc = lib.isr_inpaint_flow_static(p, p, p, 4, 512, 513, None)
    assert rc == ISR_ERR_WORKSPACE
    msg = lib.isr_last_error().decode()
    assert "ISR_INPAINT_STATIC_MAX_BYTES" in msg and str(need.value) in msg and str(MAX_BYTES) in msg
    assert lib.isr_inpaint_flow_static(None, p, p, 1, 8, 8, None) == -1       # ISR_ERR_ARG like isr_inpaint_flow
    assert lib.isr_inpaint_flow_static(p, p, p, 0, 8, 8, None) == -1


@pytest.mark.gpu
def test_static_inpaint_equals_workspace_entry_point_on_consecutive_calls():
    """Two (and more) calls in a row through the C ABI, each against isr_inpaint_flow on the same inputs: a large case, then
    a SMALLER one with different data (the scratch still holds the first call's pyramid: nothing of it may leak), the
    first again (repeatable), and the largest size the array admits (4 x 512 x 512 uses all but 16 of its bytes).  Same
    kernels, same arithmetic: the results are equal to the tolerance of the in-painting tests -- in fact bit for bit.
    The call allocates nothing: the free device memory does not shrink across it."""
    from isosurfacesuperresolution_b200 import _lib
    lib = _lib.lib()
    cases = [(2, 135, 240, 3), (1, 54, 96, 4), (2, 135, 240, 3), (4, 512, 512, 5), (1, 270, 480, 6)]
    first = None
    for k, (B, h, w, seed) in enumerate(cases):
        f_np, m_np = _case(B, h, w, seed)
        flow, hole = torch.from_numpy(f_np).cuda(), torch.from_numpy(m_np).cuda()
        ref, need = _workspace_call(lib, flow, hole)
        assert need <= MAX_BYTES
        free_before = torch.cuda.mem_get_info()[0]
        out = _static_call(lib, flow, hole)
        free_after = torch.cuda.mem_get_info()[0]
        if k > 0:        # (the very first call may load the module lazily)
            assert free_after >= free_before, (k, free_before, free_after)
        assert torch.isfinite(out).all() and torch.isfinite(ref).all()
        assert float((out - ref).abs().max()) <= TOL, (k, B, h, w)
        assert torch.equal(out, ref)
        keep = hole == 0                                            # valid pixels keep their value exactly
        assert torch.equal(out.permute(1, 0, 2, 3)[:, keep], flow.permute(1, 0, 2, 3)[:, keep])
        if k == 0:
            first = out.clone()
        if k == 2:
            assert torch.equal(out, first)
