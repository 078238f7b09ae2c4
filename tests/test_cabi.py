"""The C-ABI boundary without a GPU: libisr.so loads, exports every symbol include/isr.h declares, the ctypes
signature table covers the header, argument validation mirrors the reference's error behaviour, and compute entry
points refuse to run without a CUDA device (there is no CPU path)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "isr.h")


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(isr_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from isosurfacesuperresolution_b200 import _lib
    return _lib


def test_header_declares_expected_surface():
    names = header_functions()
    for must in ["isr_flatten_high", "isr_warp_upscale", "isr_initial_image", "isr_postprocess", "isr_shade",
                 "isr_net_create", "isr_net_load_weights", "isr_net_forward", "isr_net_forward_recurrent",
                 "isr_net_forward_recurrent_host", "isr_last_error", "isr_version"]:
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    handle = lib.lib()
    out = subprocess.check_output(["nm", "-D", "--defined-only", lib.LIB_PATH], text=True)
    exported = set(re.findall(r"\bT (isr_[a-z0-9_]+)", out))
    for name in header_functions():
        assert name in exported, name + " declared in include/isr.h but not exported"
        assert getattr(handle, name) is not None
    # no C++-mangled or torch-typed entry points in the public surface
    assert all(not n.startswith("_Z") for n in exported)


def test_ctypes_table_matches_header(lib):
    assert sorted(lib.SIGNATURES) == header_functions()


def test_version_and_error_string(lib):
    h = lib.lib()
    assert h.isr_version() >= 100
    assert isinstance(h.isr_last_error(), bytes)


def test_argument_validation_without_gpu(lib):
    h = lib.lib()
    null = ctypes.c_void_p(0)
    # flatten_high: H, W must be multiples of the factor (videotools.py:18-19 would fail in view())
    assert h.isr_flatten_high(null, null, 1, 1, 8, 8, 4, null) == lib.ERR_ARG
    # initialImage: unknown mode -> ValueError in the reference (initial_image.py:54)
    assert h.isr_initial_image(null, null, 1, 5, 4, 4, 6, 7, 0, 4, null) == lib.ERR_ARG
    assert b"unknown input mode" in h.isr_last_error()
    # shading needs >= 5 channels (shading.py:150)
    buf = (ctypes.c_float * 22)()
    assert h.isr_shade(ctypes.c_void_p(16), buf, ctypes.c_void_p(16), 1, 4, 4, 4, null) == lib.ERR_ARG
    # net: channel mask longer than the output (enhancenet.py:71-78 has no branch for it)
    net = ctypes.c_void_p(0)
    assert h.isr_net_create(ctypes.byref(net), 101, 3, 5, 1, 1) == lib.ERR_ARG
    assert h.isr_net_create(ctypes.byref(net), 101, 6, 5, 7, 1) == lib.ERR_UNSUPPORTED
    with pytest.raises(lib.IsrError):
        lib.check(h.isr_set_tunable(b"no_such_key", 1))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour WITHOUT a CUDA device")
def test_compute_entry_points_fail_loudly_without_cuda(lib):
    h = lib.lib()
    x = np.zeros((1, 1, 4, 4), np.float32)
    p = ctypes.c_void_p(x.ctypes.data)
    assert h.isr_flatten_high(p, p, 1, 1, 4, 4, 4, ctypes.c_void_p(0)) == lib.ERR_CUDA
    assert b"no CPU path" in h.isr_last_error()
    net = ctypes.c_void_p(0)
    assert h.isr_net_create(ctypes.byref(net), 101, 6, 5, 1, 1) == lib.ERR_CUDA


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour WITHOUT a CUDA device")
def test_python_api_refuses_cpu_tensors():
    import isosurfacesuperresolution_b200 as isr
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        isr.VideoTools.flatten_high(torch.zeros(1, 1, 4, 4), 4)
    from argparse import Namespace
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6,
                            Namespace(upsample="bilinear", reconType="residual", useBN=False))
    with torch.no_grad(), pytest.raises(RuntimeError, match="no CPU fallback"):
        net(torch.zeros(1, 101, 4, 4))
    with pytest.raises(ValueError):
        isr.createNetwork("nope", 4, 101, [0], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))


def test_every_tunable_is_documented_and_accepted(lib):
    """The tunables list of include/isr.h and the keys isr_api.cu understands are the same set, and every documented key is
    accepted by isr_set_tunable (no GPU needed: it only changes the defaults new handles snapshot); unknown keys and
    out-of-range values come back as a negative status, and the previous value is restored."""
    doc = open(HEADER).read()
    doc = doc[doc.index("Tunables for tests, experiments and bench ablations"):doc.index("int isr_set_tunable")]
    documented = set(re.findall(r'"([a-z_0-9]+)"', doc))
    code = open(os.path.join(ROOT, "isosurfacesuperresolution_b200", "csrc", "isr_api.cu")).read()
    understood = set(re.findall(r'strcmp\(key, "([a-z_0-9]+)"\)', code))
    assert documented == understood, (sorted(documented - understood), sorted(understood - documented))
    for key in sorted(documented):
        old = None
        for v in (1, 0, 2, 3, 10, 16):
            rc = lib.lib().isr_set_tunable(key.encode(), v)
            if rc >= 0:
                old = rc
                break
        assert old is not None, key
        assert lib.lib().isr_set_tunable(key.encode(), old) >= 0          # restore the default
        assert lib.lib().isr_set_tunable(key.encode(), old) == old
    assert lib.lib().isr_set_tunable(b"no_such_tunable", 1) < 0
    assert lib.lib().isr_set_tunable(b"trunk_full_blocks", 11) < 0
