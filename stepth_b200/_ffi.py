"""ctypes binding of libstepth_cuda.so -- exactly the symbols include/stepth_cuda.h declares.

This is the Python twin of the `stepth-cuda-sys` FFI crate (rust/stepth-cuda-sys/src/lib.rs): the same
symbols, the same argument meaning.  There is no CPU fallback: if the library is missing it is built with
nvcc, and if that is impossible the import of the product package fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _build

OK = 0
ERR_INVALID_ARG = 1
ERR_SIZE_MISMATCH = 2
ERR_DIV_BY_ZERO_MAX = 3
ERR_TOO_FEW_BLOCKS = 4
ERR_DEGENERATE_DEPTH = 5
ERR_NO_CONVERGENCE = 6
ERR_OUT_OF_BOUNDS = 7
ERR_CUDA = 100
ERR_NCCL = 101

RESAMPLE_OFF, RESAMPLE_U8, RESAMPLE_F32 = 0, 1, 2
FLAG_INVERT = 1 << 2
FLAG_FOREGROUND = 1 << 3
FLAG_APPLY = 1 << 4
FLAG_PIXELOPPS_SAT = 1 << 5
FLAG_LITERAL_SEARCH = 1 << 6

BLOCK_DTYPE = np.dtype([("x", "<u4"), ("y", "<u4"), ("w", "<u4"), ("h", "<u4"), ("value", "u1", (3,)), ("pad", "u1")])
assert BLOCK_DTYPE.itemsize == 20


class Params(C.Structure):
    _fields_ = [("precision", C.c_uint8 * 3), ("pad", C.c_uint8), ("max_rings", C.c_uint32), ("flags", C.c_uint32)]


class DevShape(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("n_pairs", "w", "h", "img_h", "y0", "wa", "ha", "add_y0", "add_img_h")]


def make_params(precision, max_rings: int = 255, flags: int = RESAMPLE_U8) -> Params:
    p = Params()
    pr = [int(v) for v in precision]
    if len(pr) != 3 or any(v < 0 or v > 255 for v in pr):
        raise ValueError("precision must be three u8 values")
    p.precision[0], p.precision[1], p.precision[2] = pr
    p.pad = 0
    p.max_rings = int(max_rings)
    p.flags = int(flags)
    return p


_vp, _u8p, _u32p, _i32p = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p  # raw addresses (host or device)
_sz, _u32, _i32, _u8 = C.c_size_t, C.c_uint32, C.c_int32, C.c_uint8
_PP = C.POINTER(Params)
_SP = C.POINTER(DevShape)

# name -> (restype, argtypes): one entry per symbol of include/stepth_cuda.h
SIGNATURES = {
    "stepth_cuda_device_count": (_i32, [C.POINTER(_i32)]),
    "stepth_cuda_ctx_create": (_i32, [_i32, C.POINTER(_vp)]),
    "stepth_cuda_ctx_destroy": (None, [_vp]),
    "stepth_cuda_last_error": (C.c_char_p, [_vp]),
    "stepth_cuda_version": (C.c_char_p, []),
    "stepth_cuda_depth_from_blocks": (_i32, [_vp, _u8p, _u32, _u32, _vp, _sz, _u32, _u32, _PP, _u8p, _u8p, _u8p, _u8p, _u8p, _u8p]),
    "stepth_cuda_depth_dense": (_i32, [_vp, _u8p, _u32, _u32, _u8p, _u32, _u32, _PP, _u8p, _u8p, _u8p, _u8p, _u8p]),
    "stepth_cuda_depth_dense_batch": (_i32, [_vp, _u32, _vp, _u32, _u32, _vp, _u32, _u32, _PP, _vp, _vp, _vp, _u8p, _i32p]),
    "stepth_cuda_depth_dense_batch_rgb": (_i32, [_vp, _u32, _vp, _u32, _u32, _vp, _u32, _u32, _PP, _vp, _vp, _u8p, _i32p]),
    "stepth_cuda_invert_depth": (_i32, [_vp, _u8p, _sz]),
    "stepth_cuda_depth_split": (_i32, [_vp, _u8p, _sz, _u8, _u8p, C.POINTER(_u32), C.POINTER(_i32)]),
    "stepth_cuda_slice": (_i32, [_vp, _u8p, _sz, _i32, _i32, _u8p]),
    "stepth_cuda_select_foreground": (_i32, [_vp, _u8p, _sz, _u8p, _u8p]),
    "stepth_cuda_apply_mask": (_i32, [_vp, _u8p, _u8p, _sz]),
    "stepth_cuda_highlight_depth": (_i32, [_vp, _u8p, _u8p, _sz, _u8p]),
    "stepth_cuda_highlight_mask": (_i32, [_vp, _u8p, _u8p, _sz, _u8p]),
    "stepth_cuda_mask_and": (_i32, [_vp, _u8p, _u8p, _sz]),
    "stepth_cuda_mask_or": (_i32, [_vp, _u8p, _u8p, _sz]),
    "stepth_cuda_mask_not": (_i32, [_vp, _u8p, _sz]),
    "stepth_cuda_image_replace": (_i32, [_vp, _u8p, _u8p, _u32, _u32, _u8p, _u32, _u32, _u32, _u32]),
    "stepth_cuda_normalize_brightness_luma16": (_i32, [_vp, _vp, _sz, _vp, _sz, C.c_double, _vp]),
    "stepth_cuda_normalize_brightness_rgb16": (_i32, [_vp, _vp, _sz, _vp, _sz, C.c_double, _vp]),
    "stepth_host_discretise_v0": (_sz, [_u8p, _u32, _u32, _u32, _u8p, _u32, _u32, _vp, _sz]),
    "stepth_cuda_discretise_v0": (_i32, [_vp, _u8p, _u32, _u32, _u32, _u8p, _u32, _u32, _vp, _sz, C.POINTER(_sz)]),
    "stepth_cuda_dev_discretise_v0": (_i32, [_vp, _vp, _u8p, _u32, _u32, _u32, _u8p, _u32, _u32, _vp, _sz, _u32p]),
    "stepth_cuda_dev_discretise_v0_noalloc": (_i32, [_vp, _vp, _u8p, _u32, _u32, _u32, _u8p, _u32, _u32, _vp, _sz, _u32p]),
    "stepth_cuda_depth_from_additional": (_i32, [_vp, _u8p, _u32, _u32, _u8p, _u32, _u32, _PP, _u32, _u32, _u8p, _u8p, _u8p, _u8p,
                                                 C.POINTER(_sz)]),
    "stepth_cuda_decode_jpeg": (_i32, [_vp, _u8p, _sz, _u32p, _u32p, _u8p, _sz]),
    "stepth_cuda_encode_jpeg": (_i32, [_vp, _u8p, _u32, _u32, _u32, _i32, _u8p, _sz, C.POINTER(_sz)]),
    "stepth_cuda_dev_search_dense": (_i32, [_vp, _vp, _SP, _PP, _u8p, _u8p, _u8p, _u32p]),
    "stepth_cuda_dev_search_blocks": (_i32, [_vp, _vp, _PP, _u8p, _u32, _u32, _vp, _sz, _u8p, _u32p]),
    "stepth_cuda_dev_paint_blocks": (_i32, [_vp, _vp, _vp, _u8p, _sz, _u32, _u32, _u8p]),
    "stepth_cuda_dev_finish": (_i32, [_vp, _vp, _u32, _u32, _u32, _u8p, _u32, _u32, _u32p, _u32, _u8p, _u32, _u32, _u32p, _i32p]),
    "stepth_cuda_dev_depth_split": (_i32, [_vp, _vp, _u32, _u32p, _u8, _u8p, _u32p, _i32p]),
    "stepth_cuda_dev_mask_apply": (_i32, [_vp, _vp, _u32, _sz, _u8p, _u8p, _u32, _i32p, _u8p, _u8p, _u8p]),
    "stepth_cuda_dev_histogram": (_i32, [_vp, _vp, _u8p, _sz, _u32p]),
    "stepth_cuda_launch_count": (C.c_uint64, [_vp]),
    "stepth_cuda_int_peak": (_i32, [_vp, _i32, C.POINTER(C.c_double)]),
    "stepth_host_band_plan": (None, [_u32, _u32, _u32, _u32, _u32, C.POINTER(_u32 * 4)]),
    "stepth_cuda_comm_unique_id": (_i32, [_u8p]),
    "stepth_cuda_comm_create": (_i32, [_vp, _i32, _i32, _u8p, C.POINTER(_vp)]),
    "stepth_cuda_comm_destroy": (None, [_vp]),
    "stepth_cuda_dev_depth_dense_band": (_i32, [_vp, _vp, _u32, _u32, _u32, _u32, _PP, _u8p, _u8p, _u8p, _u8p, _u8p, _u8p]),
    "stepth_cuda_comm_band_result": (_i32, [_vp, _vp, C.POINTER(_i32), _u8p]),
    "stepth_cuda_comm_peer_mode": (_i32, [_vp]),
    "stepth_cuda_comm_full_map": (_i32, [_vp, _vp, _u32, _u32, _u32, _u32, C.POINTER(_vp)]),
    "stepth_cuda_comm_band_stage_ms": (_i32, [_vp, C.POINTER(C.c_float * 4)]),
    "stepth_cuda_multi_create": (_i32, [C.POINTER(_i32), _u32, C.POINTER(_vp)]),
    "stepth_cuda_multi_destroy": (None, [_vp]),
    "stepth_cuda_multi_last_error": (C.c_char_p, [_vp]),
    "stepth_cuda_multi_launch_count": (C.c_uint64, [_vp]),
    "stepth_cuda_multi_depth_dense_batch": (_i32, [_vp, _u32, _vp, _u32, _u32, _vp, _u32, _u32, _PP, _vp, _vp, _vp, _u8p, _i32p]),
    "stepth_cuda_multi_depth_dense_banded": (_i32, [_vp, _u8p, _u32, _u32, _u8p, _u32, _u32, _PP, _u8p, _u8p, _u8p, _u8p]),
    "stepth_cuda_dev_count_probes": (_i32, [_vp, _vp, _SP, _PP, _u8p, _u8p, _u8p, _u32p, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Loads (building it first if it is absent) the in-tree libstepth_cuda.so."""
    global _lib
    if _lib is None:
        if not os.path.exists(_build.LIB_PATH) or os.environ.get("STEPTH_REBUILD") == "1":
            _build.build(force=True)
        # STEPTH_LIB: a tuning variant of the SAME library built by _build.build(out=..., defs=...) (kernel sweeps only)
        L = C.CDLL(os.environ.get("STEPTH_LIB") or _build.LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header and library disagree: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class StepthError(RuntimeError):
    """A non-zero status from the C ABI.  `.code` is the STEPTH_ERR_* value."""

    def __init__(self, code: int, message: str):
        super().__init__(f"[stepth status {code}] {message}")
        self.code = code


def ptr(a) -> int:
    """Address of a C-contiguous numpy array or torch tensor (host or device); None -> NULL."""
    if a is None:
        return 0
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("buffer must be C-contiguous")
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        if not a.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return a.data_ptr()
    if isinstance(a, int):
        return a
    raise TypeError(f"cannot take the address of {type(a)}")
