"""ctypes binding of libllie.so (include/llie.h).  No logic, no fallback.

The library must be the in-tree build (low_light_image_enhancement_b200/libllie.so);
a missing library raises ImportError-like RuntimeError -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# LLIE_LIB_PATH: test tooling only (tools/ovf_check.sh runs the suite on a variant build); unset = the in-tree build
LIB_PATH = os.environ.get("LLIE_LIB_PATH") or os.path.join(HERE, "libllie.so")

OK, E_ARG, E_SHAPE, E_DTYPE, E_CUDA, E_NOPEER, E_NOMEM, E_STATE = 0, -1, -2, -3, -4, -5, -6, -7
NHWC, NCHW = 0, 1
F32, BF16, F32_SIMT = 0, 1, 2
FINISH = {"none": 0, "minmax": 1, "meanstd": 2, "percentile": 3}
BLUR_PYRAMID, BLUR_DIRECT = 0, 1
METHOD = {"msrcr": 0, "zero_dce": 1}
SHARD_DEVICE, SHARD_HOST = 0, 1
MAX_SCALES = 8
STATIC_SCRATCH_BYTES, STATIC_MAX_TAPS = 32 << 20, 16384  # llie_msrcr_direct_static_f32


class LlieError(RuntimeError):
    """Any non-zero status from libllie (code in .status)."""

    def __init__(self, status: int, msg: str):
        super().__init__(f"libllie status {status}: {msg}")
        self.status = status


class LlieCudaError(LlieError):
    pass


class MsrcrParams(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("n_scales", C.c_int32),
        ("sigma", C.c_float * MAX_SCALES), ("weight", C.c_float * MAX_SCALES),
        ("truncate", C.c_float), ("alpha", C.c_float), ("beta", C.c_float), ("gain", C.c_float),
        ("offset", C.c_float), ("in_scale", C.c_float), ("in_bias", C.c_float), ("blur", C.c_int32),
        ("finish", C.c_int32), ("finish_k", C.c_float),
    ]


class DceParams(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("iters", C.c_int32), ("curve_sign", C.c_int32),
                ("precision", C.c_int32)]


# every symbol include/llie.h declares: name -> (restype, argtypes)
_VP, _I, _F = C.c_void_p, C.c_int, C.c_float
EXPORTS = {
    "llie_version": (_I, []),
    "llie_last_error": (C.c_char_p, []),
    "llie_device_count": (_I, [C.POINTER(_I)]),
    "llie_msrcr_default_params": (_I, [C.POINTER(MsrcrParams)]),
    "llie_dce_default_params": (_I, [C.POINTER(DceParams)]),
    "llie_create": (_I, [C.POINTER(_VP), _I]),
    "llie_destroy": (_I, [_VP]),
    "llie_create_group": (_I, [C.POINTER(_VP), C.POINTER(_I), _I]),
    "llie_destroy_group": (_I, [C.POINTER(_VP), _I]),
    "llie_enhance_sharded": (_I, [C.POINTER(_VP), _I, _I, _I, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_I), _I, _I, _I,
                                  _VP, _I]),
    "llie_shard_last_ms": (_I, [_VP, C.POINTER(_F)]),
    "llie_sync": (_I, [_VP]),
    "llie_workspace_bytes": (_I, [_VP, C.POINTER(C.c_size_t)]),
    "llie_launch_count": (_I, [_VP, C.POINTER(C.c_uint64)]),
    "llie_profile_begin": (_I, [_VP]),
    "llie_profile_end": (_I, [_VP, C.c_char_p, C.c_size_t]),
    "llie_msrcr_f32": (_I, [_VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(MsrcrParams), _VP]),
    "llie_msrcr_direct_static_f32": (_I, [_VP, _VP, _I, _I, _I, C.POINTER(MsrcrParams), _VP]),
    "llie_msrcr_u8": (_I, [_VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(MsrcrParams), _VP]),
    "llie_dce_forward_u8": (_I, [_VP, _VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(DceParams), _VP]),
    "llie_msrcr_u16": (_I, [_VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(MsrcrParams), _VP]),
    "llie_dce_forward_bf16io": (_I, [_VP, _VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(DceParams), _VP]),
    "llie_dce_forward_u16": (_I, [_VP, _VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(DceParams), _VP]),
    "llie_msrcr_plan_query": (_I, [_F, _F, C.POINTER(_I), C.POINTER(C.c_double), C.POINTER(_I),
                                   C.POINTER(C.c_double)]),
    "llie_msrcr_plan_errors": (_I, [_F, _F, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "llie_msrcr_plan_taps": (_I, [_F, _F, C.POINTER(_F), _I, C.POINTER(_I)]),
    "llie_dce_set_weights": (_I, [_VP, C.POINTER(_F), C.c_size_t]),
    "llie_dce_forward": (_I, [_VP, _VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(DceParams), _VP]),
    "llie_debug_dce_activation": (_I, [_VP, _I, _I, C.POINTER(_F)]),
    "llie_debug_dce_mode": (_I, [_VP, _I]),
    "llie_dce_curves": (_I, [_VP, _VP, _VP, _VP, _I, _I, _I, _I, C.POINTER(DceParams), _VP]),
    "llie_msrcr_tiled": (_I, [C.POINTER(_VP), _I, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_I), C.POINTER(_I),
                              _I, _I, C.POINTER(MsrcrParams)]),
    "llie_msrcr_tiled_last_ms": (_I, [_VP, C.POINTER(_F)]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m low_light_image_enhancement_b200.build` "
                "(there is no CPU or PyTorch fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(status: int) -> None:
    if status != OK:
        msg = load().llie_last_error().decode("utf-8", "replace")
        raise (LlieCudaError if status == E_CUDA else LlieError)(status, msg)
