"""ctypes binding of libraftcorr_b200.so (the C ABI declared in include/raftcorr.h).

There is no CPU fallback and no alternative backend: if the CUDA library is missing or a call
fails, this module raises.  Build the library with ``python __graft_entry__.py`` (or
``raftcorr_b200/build.py``).
"""
from __future__ import annotations

import ctypes
import os

PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("RAFTCORR_LIB", os.path.join(PKG, "libraftcorr_b200.so"))

MAX_LEVELS = 4
COORD_ATEN_CPU, COORD_ATEN_CUDA, COORD_DIRECT = 0, 1, 2
PREC_FP32, PREC_BF16, PREC_TF32 = 0, 1, 2
NOALLOC_SCRATCH_BYTES = 8 << 20  # RAFTCORR_NOALLOC_SCRATCH_BYTES

c_float_p = ctypes.c_void_p  # device pointers travel as integers


class Level(ctypes.Structure):
    """struct raftcorr_level"""
    _fields_ = [("data", ctypes.c_void_p), ("H", ctypes.c_int32), ("W", ctypes.c_int32),
                ("pitch", ctypes.c_int64), ("image_stride", ctypes.c_int64)]


LevelArray = Level * MAX_LEVELS


class Level16(ctypes.Structure):
    """struct raftcorr_level16: bfloat16 pixels in 4x8 blocks of 64 bytes"""
    _fields_ = [("data", ctypes.c_void_p), ("H", ctypes.c_int32), ("W", ctypes.c_int32),
                ("blocks_x", ctypes.c_int32), ("blocks_y", ctypes.c_int32),
                ("interleave_log2", ctypes.c_int32), ("reserved", ctypes.c_int32), ("image_stride", ctypes.c_int64)]


Level16Array = Level16 * MAX_LEVELS

#: every symbol include/raftcorr.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "raftcorr_version": (ctypes.c_int, []),
    "raftcorr_last_error": (ctypes.c_char_p, []),
    "raftcorr_pyramid_layout": (ctypes.c_int, [ctypes.c_int] * 4 + [ctypes.POINTER(Level),
                                                                  ctypes.POINTER(ctypes.c_int64)]),
    "raftcorr_build_workspace_bytes": (ctypes.c_int64, [ctypes.c_int] * 5),
    "raftcorr_build_pyramid": (ctypes.c_int, [c_float_p, c_float_p] + [ctypes.c_int] * 5 +
                               [ctypes.POINTER(Level), ctypes.c_int, ctypes.c_void_p, ctypes.c_int64,
                                ctypes.c_void_p]),
    "raftcorr_build_pyramid_noalloc": (ctypes.c_int, [c_float_p, c_float_p] + [ctypes.c_int] * 5 +
                                       [ctypes.POINTER(Level), ctypes.c_int, ctypes.c_void_p]),
    "raftcorr_pyramid16_layout": (ctypes.c_int, [ctypes.c_int] * 4 + [ctypes.POINTER(Level16),
                                                                    ctypes.POINTER(ctypes.c_int64)]),
    "raftcorr_build_pyramid16": (ctypes.c_int, [c_float_p, c_float_p] + [ctypes.c_int] * 5 +
                                 [ctypes.POINTER(Level16), ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "raftcorr_lookup16_fwd": (ctypes.c_int, [ctypes.POINTER(Level16), ctypes.c_int, c_float_p, c_float_p] +
                              [ctypes.c_int] * 5 + [ctypes.c_void_p]),
    "raftcorr_lookup_fwd": (ctypes.c_int, [ctypes.POINTER(Level), ctypes.c_int, c_float_p, c_float_p] +
                            [ctypes.c_int] * 5 + [ctypes.c_void_p]),
    "raftcorr_lookup_fwd_nhwc": (ctypes.c_int, [ctypes.POINTER(Level), ctypes.c_int, c_float_p, c_float_p] +
                                 [ctypes.c_int] * 5 + [ctypes.c_void_p]),
    "raftcorr_lookup_bwd": (ctypes.c_int, [ctypes.POINTER(Level), ctypes.POINTER(Level), ctypes.c_int,
                                           c_float_p, c_float_p, c_float_p] + [ctypes.c_int] * 5 +
                            [ctypes.c_void_p]),
    "raftcorr_build_bwd_workspace_bytes": (ctypes.c_int64, [ctypes.c_int] * 4),
    "raftcorr_build_bwd": (ctypes.c_int, [ctypes.POINTER(Level), ctypes.c_int, c_float_p, c_float_p,
                                          c_float_p, c_float_p] + [ctypes.c_int] * 4 +
                           [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "raftcorr_alt_fwd": (ctypes.c_int, [c_float_p] * 4 + [ctypes.c_int] * 8 + [ctypes.c_float,
                                                                              ctypes.c_void_p]),
    "raftcorr_alt_fused_fwd": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), c_float_p, c_float_p] +
                               [ctypes.c_int] * 7 + [ctypes.c_float, ctypes.c_void_p]),
    "raftcorr_alt_pack": (ctypes.c_int, [c_float_p] + [ctypes.c_int] * 5 + [ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p]),
    "raftcorr_alt_bwd": (ctypes.c_int, [c_float_p] * 7 + [ctypes.c_int] * 8 + [ctypes.c_float,
                                                                              ctypes.c_void_p]),
    "raftcorr_upsample_flow_fwd": (ctypes.c_int, [c_float_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_void_p]),
    "raftcorr_upsample_flow_bwd": (ctypes.c_int, [c_float_p] * 5 + [ctypes.c_int] * 3 + [ctypes.c_void_p]),
    "raftcorr_upflow8": (ctypes.c_int, [c_float_p] * 2 + [ctypes.c_int] * 3 + [ctypes.c_void_p]),
    "raftcorr_lookup_conv1x1_workspace_bytes": (ctypes.c_int64, [ctypes.c_int] * 5),
    "raftcorr_conv1x1_nhwc_fwd": (ctypes.c_int, [c_float_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_void_p]),
    "raftcorr_conv1x1_nchw_fwd": (ctypes.c_int, [c_float_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_void_p]),
    "raftcorr_lookup_conv1x1_fwd": (ctypes.c_int, [ctypes.POINTER(Level), ctypes.c_int] + [c_float_p] * 4 +
                                    [ctypes.c_int] * 7 + [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
}

_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA extension has not been built "
                "(run `python __graft_entry__.py` or raftcorr_b200/build.py). "
                "There is no CPU or PyTorch fallback for this path.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class RaftCorrError(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().raftcorr_last_error().decode("utf-8", "replace")
        raise RaftCorrError(f"{what} failed (code {rc}): {msg}")
