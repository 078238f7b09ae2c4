"""ctypes binding of libisr.so (include/isr.h).

Mirrors how the reference binds its own native code (inference/renderer.py:78-117: ``cdll.LoadLibrary``, declared
``argtypes`` / ``restype``, ``tensor.data_ptr()`` passed as an integer).  There is no fallback: a missing library or a
failing call raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libisr.so")

PREC = {"bf16": 0, "bf16x3": 1, "mixed": 2, "fp16": 3, "fp16x3": 4, "mixed2": 5, "mixed3": 6, "mixed4": 7, "mixed5": 8, "mixed6": 9, "hybrid": 10}
UPSAMPLE = {"nearest": 0, "bilinear": 1}
INITIAL = {"zero": 0, "unshaded": 1, "input": 2}

ERR_ARG, ERR_CUDA, ERR_WORKSPACE, ERR_UNSUPPORTED = -1, -2, -3, -4


class IsrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libisr error {code}: {msg}")
        self.code = code


_c = ctypes
_vp, _i, _sz = _c.c_void_p, _c.c_int, _c.c_size_t

# include/isr.h ISR_RC_*: float indices of the ray-caster's parameter block
(RC_CAM_POS, RC_CAM_U, RC_CAM_V, RC_CAM_S, RC_BOX_MIN, RC_VOXEL, RC_ISO, RC_STEP, RC_VIEW_CUR, RC_VIEW_NEXT, RC_NORMAL_MAT,
 RC_LIGHT, RC_AMBIENT, RC_DIFFUSE, RC_SPECULAR, RC_SPEC_EXP, RC_AO_RADIUS, RC_AO_BIAS, RC_PARAMS) = (
    0, 3, 6, 9, 12, 15, 16, 17, 18, 34, 50, 66, 69, 72, 75, 78, 79, 80, 81)

# name -> (restype, argtypes); every symbol include/isr.h declares
SIGNATURES = {
    "isr_version": (_i, []),
    "isr_last_error": (_c.c_char_p, []),
    "isr_launch_count": (_c.c_uint64, []),
    "isr_flatten_high": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "isr_warp_upscale": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "isr_initial_image": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "isr_postprocess": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "isr_shade": (_i, [_vp, _c.POINTER(_c.c_float), _vp, _i, _i, _i, _i, _vp]),
    "isr_compose": (_i, [_vp, _vp, _c.POINTER(_c.c_float), _vp, _i, _i, _i, _i, _i, _i, _i, _c.c_float, _vp]),
    "isr_ingest_gbuffer": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "isr_ingest_low_channels": (_i, [_i]),
    "isr_inpaint_workspace_bytes": (_i, [_i, _i, _i, _c.POINTER(_sz)]),
    "isr_inpaint_flow": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "isr_inpaint_flow_static": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "isr_ssao": (_i, [_i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _c.c_float, _c.c_float, _i, _vp]),
    "isr_ssao_tables": (_i, [_i, _i, _vp, _vp]),
    "isr_raycast_iso": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "isr_raycast_skip_bytes": (_i, [_i, _i, _i, _c.POINTER(_sz)]),
    "isr_raycast_build_skip": (_i, [_vp, _i, _i, _i, _vp, _sz, _vp]),
    "isr_raycast_ao_tables": (_i, [_vp, _vp]),
    "isr_net_create": (_i, [_c.POINTER(_vp), _i, _i, _i, _i, _i]),
    "isr_net_destroy": (None, [_vp]),
    "isr_net_load_weights": (_i, [_vp, _c.POINTER(_vp), _i, _vp]),
    "isr_net_workspace_bytes": (_i, [_vp, _i, _i, _i, _i, _c.POINTER(_sz)]),
    "isr_net_forward": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "isr_net_forward_recurrent": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz,
                                       _vp]),
    "isr_net_forward_recurrent_host": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _sz, _vp,
                                            _vp]),
    "isr_pipe_staging_floats": (_i, [_vp, _i, _i, _i, _i, _i, _c.POINTER(_sz)]),
    "isr_pipe_create": (_i, [_c.POINTER(_vp), _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp, _vp, _sz, _i]),
    "isr_pipe_submit": (_i, [_vp, _vp, _vp, _vp]),
    "isr_pipe_enable_display": (_i, [_vp, _c.POINTER(_c.c_float), _vp, _sz]),
    "isr_pipe_submit_display": (_i, [_vp, _vp, _vp, _vp]),
    "isr_shade_rgb8": (_i, [_vp, _c.POINTER(_c.c_float), _vp, _i, _i, _i, _i, _vp]),
    "isr_pipe_wait": (_i, [_vp]),
    "isr_pipe_destroy": (None, [_vp]),
    "isr_debug_conv3x3": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "isr_debug_conv_workspace_bytes": (_i, [_i, _i, _i, _c.POINTER(_sz)]),
    "isr_set_tunable": (_i, [_c.c_char_p, _i]),
    "isr_net_set_tunable": (_i, [_vp, _c.c_char_p, _i]),
    "isr_temporal_blend": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _c.c_double, _vp]),
    "isr_pipe_set_state": (_i, [_vp, _vp]),
    "isr_pipe_enable_f16": (_i, [_vp, _vp, _sz]),
    "isr_pipe_submit_f16": (_i, [_vp, _vp, _vp, _vp]),
    "isr_net_time_conv": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _sz, _i, _c.POINTER(_c.c_float), _vp]),
}

_LIB = None


def build(force=False):
    """Compiles libisr.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    srcs = [os.path.join(_HERE, "csrc", f) for f in os.listdir(os.path.join(_HERE, "csrc")) if not f.endswith(".o")]
    srcs.append(os.path.join(_ROOT, "include", "isr.h"))
    stale = force or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", _ROOT, "-s", os.path.relpath(LIB_PATH, _ROOT)],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU / PyTorch fallback for the inference path)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = handle
    return _LIB


def check(rc):
    if rc < 0:
        raise IsrError(rc, lib().isr_last_error().decode("utf-8", "replace"))
    return rc


def launch_count():
    return int(lib().isr_launch_count())


def set_tunable(key, value):
    rc = lib().isr_set_tunable(key.encode(), int(value))
    if rc < 0:
        raise IsrError(rc, lib().isr_last_error().decode())
    return rc
