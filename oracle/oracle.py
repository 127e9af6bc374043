"""ctypes binding of the CPU oracle (oracle/stepth_oracle.c).

TEST INFRASTRUCTURE ONLY -- PARITY UNPINNED (see stepth_oracle.h).  Import this only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  The product package
`stepth_b200` never imports it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libstepth_oracle.so")

OK, INVALID_ARG, SIZE_MISMATCH, DIV_BY_ZERO_MAX, TOO_FEW_BLOCKS, DEGENERATE_DEPTH, NO_CONVERGENCE = range(7)
PIXELOPPS_ABS_ALL, PIXELOPPS_SAT_ALL, PIXELOPPS_ABS_SUM = 0, 1, 2
RESAMPLE_OFF, RESAMPLE_U8, RESAMPLE_F32 = 0, 1, 2

BLOCK_DTYPE = np.dtype(
    [("x", "<u4"), ("y", "<u4"), ("w", "<u4"), ("h", "<u4"), ("value", "u1", (3,)), ("pad", "u1")]
)
assert BLOCK_DTYPE.itemsize == 20


class OracleError(RuntimeError):
    def __init__(self, code: int):
        super().__init__(f"oracle status {code}")
        self.code = code


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "stepth_oracle.c")
    hdr = os.path.join(_HERE, "stepth_oracle.h")
    stale = (
        force
        or not os.path.exists(_LIB_PATH)
        or (os.path.exists(src) and os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr)))
    )
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
        L.orc_distance_dot_dot.restype = C.c_uint32
        L.orc_distance_dot_dot.argtypes = [C.c_uint32] * 4
        L.orc_distance_dot_array.restype = C.c_int
        L.orc_distance_dot_array.argtypes = [u8p, u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                             u8p, C.c_int, u32p, u32p, u32p, u64p]
        L.orc_search_blocks.restype = C.c_int32
        L.orc_search_blocks.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, u8p, C.c_uint32,
                                        C.c_int, C.c_int, u8p, u64p]
        L.orc_search_dense.restype = C.c_int32
        L.orc_search_dense.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p, C.c_uint32, C.c_uint32, u8p,
                                       C.c_uint32, C.c_int, C.c_int, u8p, u64p]
        L.orc_search_dense_subset.restype = C.c_int32
        L.orc_search_dense_subset.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p, C.c_uint32, C.c_uint32,
                                              u8p, C.c_uint32, C.c_int, C.c_int, u32p, C.c_size_t, u8p, u64p]
        L.orc_normalise.restype = C.c_int32
        L.orc_normalise.argtypes = [u8p, C.c_size_t, u8p]
        L.orc_paint_blocks.restype = C.c_int32
        L.orc_paint_blocks.argtypes = [C.c_void_p, u8p, C.c_size_t, C.c_uint32, C.c_uint32, u8p]
        L.orc_gaussian_resample_unit.restype = C.c_int32
        L.orc_gaussian_resample_unit.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_int, u8p]
        L.orc_depth_from_blocks.restype = C.c_int32
        L.orc_depth_from_blocks.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, u8p, C.c_uint32,
                                            C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, u8p, u8p]
        L.orc_depth_dense.restype = C.c_int32
        L.orc_depth_dense.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p, C.c_uint32, C.c_uint32, u8p,
                                      C.c_uint32, C.c_int, C.c_int, C.c_int, u8p, u8p]
        L.orc_invert_depth.restype = None
        L.orc_invert_depth.argtypes = [u8p, C.c_size_t]
        L.orc_depth_split.restype = C.c_int32
        L.orc_depth_split.argtypes = [u8p, C.c_size_t, C.c_uint8, u8p, u32p, C.POINTER(C.c_int), u8p, u32p]
        L.orc_slice.restype = None
        L.orc_slice.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p]
        L.orc_select_foreground.restype = C.c_int32
        L.orc_select_foreground.argtypes = [u8p, C.c_size_t, u8p, u8p]
        L.orc_apply_mask.restype = None
        L.orc_apply_mask.argtypes = [u8p, u8p, C.c_size_t]
        u16p = C.POINTER(C.c_uint16)
        L.orc_normalize_brightness_luma16.restype = None
        L.orc_normalize_brightness_luma16.argtypes = [u16p, C.c_size_t, u16p, C.c_size_t, C.c_double, u16p]
        L.orc_normalize_brightness_rgb16.restype = None
        L.orc_normalize_brightness_rgb16.argtypes = [u16p, C.c_size_t, u16p, C.c_size_t, C.c_double, u16p]
        _lib = L
    return _lib


def _u8(a: np.ndarray):
    assert a.dtype == np.uint8 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _prec(p) -> np.ndarray:
    p = np.ascontiguousarray(np.asarray(p, dtype=np.uint8).reshape(3))
    return p


def _check(st: int):
    if st != OK:
        raise OracleError(st)


def distance_dot_dot(x1, y1, x2, y2) -> int:
    return int(lib().orc_distance_dot_dot(x1, y1, x2, y2))


def distance_dot_array(what, add_rgb: np.ndarray, from_xy, max_rings: int, precision, pixelopps=PIXELOPPS_ABS_ALL):
    """helpers.rs:9-54 -> None | (distance, (x, y), probes)"""
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    ha, wa = add_rgb.shape[:2]
    w3 = _prec(what)
    p3 = _prec(precision)
    d, px, py, pr = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), C.c_uint64(0)
    hit = lib().orc_distance_dot_array(_u8(w3), _u8(add_rgb), wa, ha, int(from_xy[0]), int(from_xy[1]), int(max_rings),
                                       _u8(p3), pixelopps, C.byref(d), C.byref(px), C.byref(py), C.byref(pr))
    if not hit:
        return None
    return int(d.value), (int(px.value), int(py.value)), int(pr.value)


def search_blocks(add_rgb, blocks, precision, max_rings=255, pixelopps=PIXELOPPS_ABS_ALL, threads=0, want_probes=False):
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    blocks = np.ascontiguousarray(blocks, dtype=BLOCK_DTYPE)
    ha, wa = add_rgb.shape[:2]
    out = np.zeros(len(blocks), dtype=np.uint8)
    p3 = _prec(precision)
    pr = C.c_uint64(0)
    _check(lib().orc_search_blocks(_u8(add_rgb), wa, ha, blocks.ctypes.data, len(blocks), _u8(p3), max_rings, pixelopps,
                                   threads, _u8(out), C.byref(pr)))
    return (out, int(pr.value)) if want_probes else out


def search_dense(main_px, add_rgb, precision, max_rings=255, pixelopps=PIXELOPPS_ABS_ALL, threads=0, want_probes=False):
    main_px = np.ascontiguousarray(main_px, dtype=np.uint8)
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    h, w, stride = main_px.shape
    ha, wa = add_rgb.shape[:2]
    out = np.zeros((h, w), dtype=np.uint8)
    p3 = _prec(precision)
    pr = C.c_uint64(0)
    _check(lib().orc_search_dense(_u8(main_px), stride, w, h, _u8(add_rgb), wa, ha, _u8(p3), max_rings, pixelopps,
                                  threads, _u8(out), C.byref(pr)))
    return (out, int(pr.value)) if want_probes else out


def last_unique_probes() -> int:
    """in-bounds probes of the last search_* call, every pixel of a ring counted once (SURVEY.md 8(d) work figure)"""
    f = lib().orc_last_unique_probes
    f.restype = C.c_uint64
    f.argtypes = []
    return int(f())


def search_dense_subset(main_px, add_rgb, precision, indices, max_rings=255, pixelopps=PIXELOPPS_ABS_ALL, threads=0):
    main_px = np.ascontiguousarray(main_px, dtype=np.uint8)
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    h, w, stride = main_px.shape
    ha, wa = add_rgb.shape[:2]
    out = np.zeros(len(indices), dtype=np.uint8)
    p3 = _prec(precision)
    pr = C.c_uint64(0)
    _check(lib().orc_search_dense_subset(_u8(main_px), stride, w, h, _u8(add_rgb), wa, ha, _u8(p3), max_rings,
                                         pixelopps, threads, indices.ctypes.data_as(C.POINTER(C.c_uint32)),
                                         len(indices), _u8(out), C.byref(pr)))
    return out, int(pr.value)


def normalise(d8: np.ndarray):
    out = np.ascontiguousarray(d8, dtype=np.uint8).copy()
    mx = C.c_uint8(0)
    _check(lib().orc_normalise(_u8(out), out.size, C.byref(mx)))
    return out, int(mx.value)


def paint_blocks(blocks, values, w, h):
    blocks = np.ascontiguousarray(blocks, dtype=BLOCK_DTYPE)
    values = np.ascontiguousarray(values, dtype=np.uint8)
    out = np.zeros((h, w), dtype=np.uint8)
    _check(lib().orc_paint_blocks(blocks.ctypes.data, _u8(values), len(blocks), w, h, _u8(out)))
    return out


def gaussian_resample_unit(src: np.ndarray, mode=RESAMPLE_U8):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    h, w = src.shape
    out = np.zeros_like(src)
    _check(lib().orc_gaussian_resample_unit(_u8(src), w, h, mode, _u8(out)))
    return out


def depth_from_blocks(add_rgb, blocks, precision, w, h, max_rings=255, pixelopps=PIXELOPPS_ABS_ALL,
                      resample=RESAMPLE_U8, threads=0, want_raw=False):
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    blocks = np.ascontiguousarray(blocks, dtype=BLOCK_DTYPE)
    ha, wa = add_rgb.shape[:2]
    out = np.zeros((h, w), dtype=np.uint8)
    raw = np.zeros(len(blocks), dtype=np.uint8)
    p3 = _prec(precision)
    _check(lib().orc_depth_from_blocks(_u8(add_rgb), wa, ha, blocks.ctypes.data, len(blocks), _u8(p3), max_rings, w, h,
                                       pixelopps, resample, threads, _u8(out), _u8(raw)))
    return (out, raw) if want_raw else out


def depth_dense(main_px, add_rgb, precision, max_rings=255, pixelopps=PIXELOPPS_ABS_ALL, resample=RESAMPLE_U8,
                threads=0, want_raw=False):
    main_px = np.ascontiguousarray(main_px, dtype=np.uint8)
    add_rgb = np.ascontiguousarray(add_rgb, dtype=np.uint8)
    h, w, stride = main_px.shape
    ha, wa = add_rgb.shape[:2]
    out = np.zeros((h, w), dtype=np.uint8)
    raw = np.zeros((h, w), dtype=np.uint8)
    p3 = _prec(precision)
    _check(lib().orc_depth_dense(_u8(main_px), stride, w, h, _u8(add_rgb), wa, ha, _u8(p3), max_rings, pixelopps,
                                 resample, threads, _u8(out), _u8(raw)))
    return (out, raw) if want_raw else out


def invert_depth(depth: np.ndarray):
    out = np.ascontiguousarray(depth, dtype=np.uint8).copy()
    lib().orc_invert_depth(_u8(out), out.size)
    return out


def depth_split(depth: np.ndarray, zones: int, want_detail=False):
    """depth_image.rs:162-218 -> list of (min, max) | [(None, None)]"""
    depth = np.ascontiguousarray(depth, dtype=np.uint8)
    ranges = np.zeros(512, dtype=np.uint8)
    cents = np.zeros(256, dtype=np.uint8)
    n_out, has, iters = C.c_uint32(0), C.c_int(0), C.c_uint32(0)
    _check(lib().orc_depth_split(_u8(depth), depth.size, zones, _u8(ranges), C.byref(n_out), C.byref(has), _u8(cents),
                                 C.byref(iters)))
    if not has.value:
        res = [(None, None)]
    else:
        res = [(int(ranges[2 * i]), int(ranges[2 * i + 1])) for i in range(n_out.value)]
    if want_detail:
        return res, [int(c) for c in cents[: n_out.value]], int(iters.value)
    return res


def slice_mask(depth: np.ndarray, frm, to):
    depth = np.ascontiguousarray(depth, dtype=np.uint8)
    out = np.zeros_like(depth)
    lib().orc_slice(_u8(depth), depth.size, -1 if frm is None else int(frm), -1 if to is None else int(to), _u8(out))
    return out


def select_foreground(depth: np.ndarray):
    depth = np.ascontiguousarray(depth, dtype=np.uint8)
    out = np.zeros_like(depth)
    ft = np.zeros(2, dtype=np.uint8)
    _check(lib().orc_select_foreground(_u8(depth), depth.size, _u8(out), _u8(ft)))
    return out, (int(ft[0]), int(ft[1]))


def apply_mask(rgba: np.ndarray, mask: np.ndarray):
    out = np.ascontiguousarray(rgba, dtype=np.uint8).copy()
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    lib().orc_apply_mask(_u8(out), _u8(mask), mask.size)
    return out


def normalize_brightness_luma16(img1, img2, percent):
    img1 = np.ascontiguousarray(img1, dtype=np.uint16)
    img2 = np.ascontiguousarray(img2, dtype=np.uint16)
    out = np.zeros_like(img1)
    p = C.POINTER(C.c_uint16)
    lib().orc_normalize_brightness_luma16(img1.ctypes.data_as(p), img1.size, img2.ctypes.data_as(p), img2.size,
                                          float(percent), out.ctypes.data_as(p))
    return out


def normalize_brightness_rgb16(img1, img2, percent):
    img1 = np.ascontiguousarray(img1, dtype=np.uint16)
    img2 = np.ascontiguousarray(img2, dtype=np.uint16)
    out = np.zeros_like(img1)
    p = C.POINTER(C.c_uint16)
    lib().orc_normalize_brightness_rgb16(img1.ctypes.data_as(p), img1.size // 3, img2.ctypes.data_as(p),
                                         img2.size // 3, float(percent), out.ctypes.data_as(p))
    return out


# ---- editor operations ---------------------------------------------------------------------------
def _bind_edit():
    L = lib()
    if getattr(L, "_edit_bound", False):
        return L
    u8p = C.POINTER(C.c_uint8)
    L.orc_highlight_depth.restype = None
    L.orc_highlight_depth.argtypes = [u8p, u8p, C.c_size_t, u8p]
    L.orc_highlight_mask.restype = None
    L.orc_highlight_mask.argtypes = [u8p, u8p, C.c_size_t, u8p]
    for name in ("orc_mask_and", "orc_mask_or"):
        getattr(L, name).restype = None
        getattr(L, name).argtypes = [u8p, u8p, C.c_size_t]
    L.orc_mask_not.restype = None
    L.orc_mask_not.argtypes = [u8p, C.c_size_t]
    L.orc_image_replace.restype = C.c_int32
    L.orc_image_replace.argtypes = [u8p, u8p, C.c_uint32, C.c_uint32, u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    L._edit_bound = True
    return L


def highlight_depth(rgba: np.ndarray, depth: np.ndarray) -> np.ndarray:
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    depth = np.ascontiguousarray(depth, dtype=np.uint8)
    out = np.empty_like(rgba)
    _bind_edit().orc_highlight_depth(_u8(rgba), _u8(depth), depth.size, _u8(out))
    return out


def highlight_mask(rgba: np.ndarray, mask: np.ndarray) -> np.ndarray:
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    out = np.empty_like(rgba)
    _bind_edit().orc_highlight_mask(_u8(rgba), _u8(mask), mask.size, _u8(out))
    return out


def mask_logic(op: str, mask: np.ndarray, other: np.ndarray | None = None) -> np.ndarray:
    out = np.ascontiguousarray(mask, dtype=np.uint8).copy()
    L = _bind_edit()
    if op == "not":
        L.orc_mask_not(_u8(out), out.size)
    else:
        other = np.ascontiguousarray(other, dtype=np.uint8)
        (L.orc_mask_and if op == "and" else L.orc_mask_or)(_u8(out), _u8(other), out.size)
    return out


def image_replace(rgba: np.ndarray, mask: np.ndarray, other_rgba: np.ndarray, start_x: int, start_y: int) -> np.ndarray:
    out = np.ascontiguousarray(rgba, dtype=np.uint8).copy()
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    other_rgba = np.ascontiguousarray(other_rgba, dtype=np.uint8)
    h, w = out.shape[:2]
    oh, ow = other_rgba.shape[:2]
    _check(_bind_edit().orc_image_replace(_u8(out), _u8(mask), w, h, _u8(other_rgba), ow, oh, start_x, start_y))
    return out
