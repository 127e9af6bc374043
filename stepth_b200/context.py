"""The library context (one CUDA device + stream + scratch) and the host-buffer calls of the C ABI.

Every method here is a thin, typed wrapper over one `stepth_cuda_*` symbol: numpy in, numpy out, status codes
turned into exceptions that mirror the reference's error behaviour:

  * `StepthIOError`  <- io::Error(InvalidInput, ..)   depth_image.rs:16,47,70,85
  * `StepthPanic`    <- the arithmetic panics of the reference (divide by zero depth_image.rs:127,
                        par_chunks_mut(0) :111-112, step_by(0)/underflow/unwrap(None) :212-216)
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

from . import _ffi
from ._ffi import BLOCK_DTYPE, StepthError


class StepthIOError(OSError):
    """io::Error(ErrorKind::InvalidInput, msg) of the reference."""

    kind = "InvalidInput"


class StepthPanic(RuntimeError):
    """A Rust panic of the reference, surfaced as an exception; `.code` is the C-ABI status."""

    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


_PANIC_TEXT = {
    _ffi.ERR_DIV_BY_ZERO_MAX: "attempt to divide by zero",
    _ffi.ERR_TOO_FEW_BLOCKS: "chunk_size must not be zero",
    _ffi.ERR_DEGENERATE_DEPTH: "depth_split on a degenerate depth range (step_by(0) / subtract with overflow / unwrap on None)",
    _ffi.ERR_NO_CONVERGENCE: "depth_split did not converge (the reference recurses without bound)",
    _ffi.ERR_OUT_OF_BOUNDS: "Image index out of bounds",
}


def _u8c(a, shape_tail=None) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if shape_tail is not None and a.shape[-len(shape_tail):] != tuple(shape_tail):
        raise StepthIOError(f"expected trailing shape {shape_tail}, got {a.shape}")
    return a


class Context:
    """`stepth_ctx`: one device, internally serialised.  Use one Context per thread / per GPU."""

    def __init__(self, device: int | None = None):
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        self._lib = _ffi.lib()
        h = C.c_void_p()
        rc = self._lib.stepth_cuda_ctx_create(int(device), C.byref(h))
        if rc != _ffi.OK:
            msg = self._lib.stepth_cuda_last_error(None).decode()
            raise StepthError(rc, msg + " -- stepth_b200 has no CPU fallback")
        self._h = h
        self.device = int(device)
        self._lock = threading.Lock()

    # ---- plumbing --------------------------------------------------------------------------------------
    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.stepth_cuda_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_error(self) -> str:
        return self._lib.stepth_cuda_last_error(self._h).decode()

    def launch_count(self) -> int:
        return int(self._lib.stepth_cuda_launch_count(self._h))

    def int_peak(self, logic_only: bool = False) -> float:
        """measured 32-bit logic/add ops per second of this device (measurement aid for bench.py's int roofline)"""
        import ctypes as C

        v = C.c_double(0.0)
        self.check(self._lib.stepth_cuda_int_peak(self._h, int(logic_only), C.byref(v)))
        return float(v.value)

    def check(self, rc: int):
        if rc == _ffi.OK:
            return
        if rc in _PANIC_TEXT:
            raise StepthPanic(rc, _PANIC_TEXT[rc])
        if rc == _ffi.ERR_SIZE_MISMATCH:
            raise StepthIOError("Sizes don't match")
        if rc == _ffi.ERR_INVALID_ARG:
            raise StepthIOError(self.last_error() or "invalid argument")
        raise StepthError(rc, self.last_error())

    # ---- depth_image.rs:110-134 (+ fused tail) -----------------------------------------------------------
    def depth_dense(self, main_rgba, add_rgb, precision, max_rings=255, flags=_ffi.RESAMPLE_U8, want_raw=False):
        """Every main pixel is its own 1x1 block.  Returns a dict: depth [, raw, mask, rgba, from_to]."""
        main_rgba = _u8c(main_rgba, (4,))
        add_rgb = _u8c(add_rgb, (3,))
        h, w = main_rgba.shape[:2]
        ha, wa = add_rgb.shape[:2]
        p = _ffi.make_params(precision, max_rings, flags)
        out = {"depth": np.empty((h, w), np.uint8)}
        raw = np.empty((h, w), np.uint8) if want_raw else None
        mask = np.empty((h, w), np.uint8) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = np.empty((h, w, 4), np.uint8) if flags & _ffi.FLAG_APPLY else None
        ft = np.zeros(2, np.uint8)
        with self._lock:
            rc = self._lib.stepth_cuda_depth_dense(self._h, _ffi.ptr(main_rgba), w, h, _ffi.ptr(add_rgb), wa, ha,
                                                   C.byref(p), _ffi.ptr(out["depth"]), _ffi.ptr(raw), _ffi.ptr(rgba),
                                                   _ffi.ptr(mask), _ffi.ptr(ft))
            if raw is not None and rc in (_ffi.OK, _ffi.ERR_DIV_BY_ZERO_MAX):
                out["raw"] = raw
            try:
                self.check(rc)
            except StepthPanic as e:
                e.partial = out
                raise
        if mask is not None:
            out["mask"] = mask
            out["from_to"] = (int(ft[0]), int(ft[1]))
        if rgba is not None:
            out["rgba"] = rgba
        return out

    def discretise_v0(self, main_px, precision, min_splits: int = 16, max_splits: int | None = None) -> np.ndarray:
        """The stand-in discretiser on the GPU (csrc/disc.cuh): the same block list as the host `discretise_v0`, order included."""
        main_px = _u8c(main_px)
        h, w, stride = main_px.shape
        if max_splits is None:
            max_splits = _max_splits(w, h)
        prec = np.ascontiguousarray(np.asarray(precision, dtype=np.uint8).reshape(3))
        n = C.c_size_t(0)
        with self._lock:
            self.check(self._lib.stepth_cuda_discretise_v0(self._h, _ffi.ptr(main_px), stride, w, h, _ffi.ptr(prec), min_splits,
                                                           max_splits, None, 0, C.byref(n)))
            blocks = np.zeros(n.value, dtype=BLOCK_DTYPE)
            if n.value:
                self.check(self._lib.stepth_cuda_discretise_v0(self._h, _ffi.ptr(main_px), stride, w, h, _ffi.ptr(prec), min_splits,
                                                               max_splits, blocks.ctypes.data, n.value, C.byref(n)))
        return blocks

    def depth_from_additional(self, main_rgba, add_rgb, precision, max_rings=255, flags=_ffi.RESAMPLE_U8, min_splits: int = 16,
                              max_splits: int | None = None):
        """load_depth_from_additional in one call (depth_image.rs:91-136): the stand-in discretiser runs on the GPU and its block
        list stays there.  Returns a dict: depth, n_blocks [, mask, rgba, from_to]."""
        main_rgba = _u8c(main_rgba, (4,))
        add_rgb = _u8c(add_rgb, (3,))
        h, w = main_rgba.shape[:2]
        ha, wa = add_rgb.shape[:2]
        if max_splits is None:
            max_splits = _max_splits(w, h)
        p = _ffi.make_params(precision, max_rings, flags)
        out = {"depth": np.empty((h, w), np.uint8)}
        mask = np.empty((h, w), np.uint8) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = np.empty((h, w, 4), np.uint8) if flags & _ffi.FLAG_APPLY else None
        ft = np.zeros(2, np.uint8)
        n = C.c_size_t(0)
        with self._lock:
            self.check(self._lib.stepth_cuda_depth_from_additional(self._h, _ffi.ptr(main_rgba), w, h, _ffi.ptr(add_rgb), wa, ha,
                                                                   C.byref(p), min_splits, max_splits, _ffi.ptr(out["depth"]),
                                                                   _ffi.ptr(rgba), _ffi.ptr(mask), _ffi.ptr(ft), C.byref(n)))
        out["n_blocks"] = int(n.value)
        if mask is not None:
            out["mask"], out["from_to"] = mask, (int(ft[0]), int(ft[1]))
        if rgba is not None:
            out["rgba"] = rgba
        return out

    # ---- decode / encode on the device (csrc/codec.cuh: nvJPEG; NOT on the parity path) ---------------------------------------
    def decode_jpeg(self, data: bytes) -> np.ndarray:
        """JPEG bytes -> RGB8 [h, w, 3], decoded on the GPU.  Differs from the reference's jpeg-decoder by a level or two, as any two
        JPEG decoders do; raises StepthIOError("Failed to open image") for a stream that is not a JPEG."""
        buf = np.frombuffer(data, dtype=np.uint8)
        w, h = C.c_uint32(0), C.c_uint32(0)
        with self._lock:
            self.check(self._lib.stepth_cuda_decode_jpeg(self._h, _ffi.ptr(buf), buf.size, C.byref(w), C.byref(h), None, 0))
            out = np.empty((h.value, w.value, 3), np.uint8)
            self.check(self._lib.stepth_cuda_decode_jpeg(self._h, _ffi.ptr(buf), buf.size, C.byref(w), C.byref(h), _ffi.ptr(out), out.size))
        return out

    def encode_jpeg(self, pixels, quality: int = 75) -> bytes:
        """Luma8 [h, w], RGB8 [h, w, 3] or RGBA8 [h, w, 4] (alpha dropped) -> JPEG bytes, encoded on the GPU"""
        px = np.ascontiguousarray(pixels, dtype=np.uint8)
        ch = 1 if px.ndim == 2 else int(px.shape[2])
        h, w = px.shape[:2]
        n = C.c_size_t(0)
        with self._lock:
            self.check(self._lib.stepth_cuda_encode_jpeg(self._h, _ffi.ptr(px), w, h, ch, int(quality), None, 0, C.byref(n)))
            out = np.empty(n.value, np.uint8)
            self.check(self._lib.stepth_cuda_encode_jpeg(self._h, _ffi.ptr(px), w, h, ch, int(quality), _ffi.ptr(out), out.size, C.byref(n)))
        return out[: n.value].tobytes()

    def depth_from_blocks(self, add_rgb, blocks, w, h, precision, max_rings=255, flags=_ffi.RESAMPLE_U8,
                          rgba_in=None, want_raw=False):
        """depth_image.rs:110-134 from a DiscretePixel list (the discretiser stays on the host)."""
        add_rgb = _u8c(add_rgb, (3,))
        blocks = np.ascontiguousarray(blocks, dtype=BLOCK_DTYPE)
        ha, wa = add_rgb.shape[:2]
        p = _ffi.make_params(precision, max_rings, flags)
        out = {"depth": np.empty((h, w), np.uint8)}
        raw = np.empty(len(blocks), np.uint8) if want_raw else None
        mask = np.empty((h, w), np.uint8) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = None
        if flags & _ffi.FLAG_APPLY:
            rgba_in = _u8c(rgba_in, (4,))
            if rgba_in.shape[:2] != (h, w):
                raise StepthIOError("Sizes don't match")
            rgba = np.empty((h, w, 4), np.uint8)
        ft = np.zeros(2, np.uint8)
        with self._lock:
            rc = self._lib.stepth_cuda_depth_from_blocks(self._h, _ffi.ptr(add_rgb), wa, ha, blocks.ctypes.data,
                                                         len(blocks), w, h, C.byref(p), _ffi.ptr(out["depth"]),
                                                         _ffi.ptr(raw), _ffi.ptr(rgba_in if rgba is not None else None),
                                                         _ffi.ptr(rgba), _ffi.ptr(mask), _ffi.ptr(ft))
            self.check(rc)
        if raw is not None:
            out["raw"] = raw
        if mask is not None:
            out["mask"] = mask
            out["from_to"] = (int(ft[0]), int(ft[1]))
        if rgba is not None:
            out["rgba"] = rgba
        return out

    def depth_dense_batch(self, mains, adds, precision, max_rings=255, flags=_ffi.RESAMPLE_U8, out=None):
        """A batch of same-shape pairs.  `mains`/`adds` are sequences of host arrays (numpy or pinned torch); the main
        images are RGBA8 [h,w,4] or -- for calls without FLAG_APPLY -- RGB8 [h,w,3] (depth_image.rs:105 holds RGB).
        Returns dict(depth=[...], mask=[...], rgba=[...], from_to=ndarray[n,2], status=ndarray[n])."""
        n = len(mains)
        if n == 0 or len(adds) != n:
            raise StepthIOError("Sizes don't match")
        h, w = mains[0].shape[:2]
        ha, wa = adds[0].shape[:2]
        stride = int(mains[0].shape[2]) if len(mains[0].shape) == 3 else 0
        if stride not in (3, 4):
            raise StepthIOError("main images must be RGB8 or RGBA8")
        for m, a in zip(mains, adds):
            if tuple(m.shape) != (h, w, stride) or tuple(a.shape) != (ha, wa, 3):
                raise StepthIOError("Sizes don't match")
        if stride == 3 and flags & _ffi.FLAG_APPLY:
            raise StepthIOError("FLAG_APPLY returns RGBA pixels: pass RGBA main images")
        p = _ffi.make_params(precision, max_rings, flags)
        out = out or {}
        depth = out.get("depth") or [np.empty((h, w), np.uint8) for _ in range(n)]
        mask = (out.get("mask") or [np.empty((h, w), np.uint8) for _ in range(n)]) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = (out.get("rgba") or [np.empty((h, w, 4), np.uint8) for _ in range(n)]) if flags & _ffi.FLAG_APPLY else None
        ft = np.zeros((n, 2), np.uint8)
        status = np.zeros(n, np.int32)

        def parr(seq):
            if seq is None:
                return None
            return (C.c_void_p * n)(*[_ffi.ptr(x) for x in seq])

        pm, pa, pd, pr, pk = parr(mains), parr(adds), parr(depth), parr(rgba), parr(mask)
        with self._lock:
            if stride == 3:
                rc = self._lib.stepth_cuda_depth_dense_batch_rgb(self._h, n, pm, w, h, pa, wa, ha, C.byref(p), pd, pk,
                                                                 _ffi.ptr(ft), _ffi.ptr(status))
            else:
                rc = self._lib.stepth_cuda_depth_dense_batch(self._h, n, pm, w, h, pa, wa, ha, C.byref(p), pd, pr, pk,
                                                             _ffi.ptr(ft), _ffi.ptr(status))
            self.check(rc)
        return {"depth": depth, "mask": mask, "rgba": rgba, "from_to": ft, "status": status}

    # ---- tail ------------------------------------------------------------------------------------------------
    def invert_depth(self, depth) -> np.ndarray:
        out = _u8c(depth).copy()
        with self._lock:
            self.check(self._lib.stepth_cuda_invert_depth(self._h, _ffi.ptr(out), out.size))
        return out

    def depth_split(self, depth, zones: int):
        depth = _u8c(depth)
        ranges = np.zeros(512, np.uint8)
        n_out, has = C.c_uint32(0), C.c_int32(0)
        with self._lock:
            self.check(self._lib.stepth_cuda_depth_split(self._h, _ffi.ptr(depth), depth.size, int(zones),
                                                         _ffi.ptr(ranges), C.byref(n_out), C.byref(has)))
        if not has.value:
            return [(None, None)]
        return [(int(ranges[2 * i]), int(ranges[2 * i + 1])) for i in range(n_out.value)]

    def slice(self, depth, frm, to) -> np.ndarray:
        depth = _u8c(depth)
        mask = np.empty_like(depth)
        with self._lock:
            self.check(self._lib.stepth_cuda_slice(self._h, _ffi.ptr(depth), depth.size, -1 if frm is None else int(frm),
                                                   -1 if to is None else int(to), _ffi.ptr(mask)))
        return mask

    def select_foreground(self, depth):
        depth = _u8c(depth)
        mask = np.empty_like(depth)
        ft = np.zeros(2, np.uint8)
        with self._lock:
            self.check(self._lib.stepth_cuda_select_foreground(self._h, _ffi.ptr(depth), depth.size, _ffi.ptr(mask),
                                                               _ffi.ptr(ft)))
        return mask, (int(ft[0]), int(ft[1]))

    def apply_mask(self, rgba, mask) -> np.ndarray:
        out = _u8c(rgba, (4,)).copy()
        mask = _u8c(mask)
        if mask.size * 4 != out.size:
            raise StepthIOError("Sizes don't match")
        with self._lock:
            self.check(self._lib.stepth_cuda_apply_mask(self._h, _ffi.ptr(out), _ffi.ptr(mask), mask.size))
        return out

    # ---- editor operations ---------------------------------------------------------------------
    def highlight_depth(self, rgba, depth) -> np.ndarray:  # depth_image.rs:51-63
        rgba, depth = _u8c(rgba, (4,)), _u8c(depth)
        if depth.size * 4 != rgba.size:
            raise StepthIOError("Sizes don't match")
        out = np.empty_like(rgba)
        with self._lock:
            self.check(self._lib.stepth_cuda_highlight_depth(self._h, _ffi.ptr(rgba), _ffi.ptr(depth), depth.size, _ffi.ptr(out)))
        return out

    def highlight_mask(self, rgba, mask) -> np.ndarray:  # mask_image.rs:57-73
        rgba, mask = _u8c(rgba, (4,)), _u8c(mask)
        if mask.size * 4 != rgba.size:
            raise StepthIOError("Sizes don't match")
        out = np.empty_like(rgba)
        with self._lock:
            self.check(self._lib.stepth_cuda_highlight_mask(self._h, _ffi.ptr(rgba), _ffi.ptr(mask), mask.size, _ffi.ptr(out)))
        return out

    def mask_logic(self, op: str, mask, other=None) -> np.ndarray:  # mask_image.rs:147-197
        out = _u8c(mask).copy()
        with self._lock:
            if op == "not":
                self.check(self._lib.stepth_cuda_mask_not(self._h, _ffi.ptr(out), out.size))
            else:
                other = _u8c(other)
                if other.shape != out.shape:
                    raise StepthIOError("Sizes don't match")  # the reference would Gaussian-resize `other`: out of scope
                fn = self._lib.stepth_cuda_mask_and if op == "and" else self._lib.stepth_cuda_mask_or
                self.check(fn(self._h, _ffi.ptr(out), _ffi.ptr(other), out.size))
        return out

    def image_replace(self, rgba, mask, other_rgba, start_x: int, start_y: int) -> np.ndarray:  # mask_image.rs:99-109
        out = _u8c(rgba, (4,)).copy()
        mask, other_rgba = _u8c(mask), _u8c(other_rgba, (4,))
        h, w = out.shape[:2]
        oh, ow = other_rgba.shape[:2]
        if mask.shape != (h, w):
            raise StepthIOError("Sizes don't match")
        with self._lock:
            self.check(self._lib.stepth_cuda_image_replace(self._h, _ffi.ptr(out), _ffi.ptr(mask), w, h, _ffi.ptr(other_rgba),
                                                           ow, oh, int(start_x), int(start_y)))
        return out

    # ---- operations.rs ------------------------------------------------------------------------------------
    def normalize_brightness_luma16(self, img1, img2, percent: float) -> np.ndarray:
        img1 = np.ascontiguousarray(img1, dtype=np.uint16)
        img2 = np.ascontiguousarray(img2, dtype=np.uint16)
        out = np.empty_like(img1)
        with self._lock:
            self.check(self._lib.stepth_cuda_normalize_brightness_luma16(self._h, _ffi.ptr(img1), img1.size, _ffi.ptr(img2),
                                                                         img2.size, float(percent), _ffi.ptr(out)))
        return out

    def normalize_brightness_rgb16(self, img1, img2, percent: float) -> np.ndarray:
        img1 = np.ascontiguousarray(img1, dtype=np.uint16)
        img2 = np.ascontiguousarray(img2, dtype=np.uint16)
        if img1.size % 3 or img2.size % 3:
            raise StepthIOError("RGB16 buffers must hold whole pixels")
        out = np.empty_like(img1)
        with self._lock:
            self.check(self._lib.stepth_cuda_normalize_brightness_rgb16(self._h, _ffi.ptr(img1), img1.size // 3,
                                                                        _ffi.ptr(img2), img2.size // 3, float(percent),
                                                                        _ffi.ptr(out)))
        return out


class Multi:
    """One process driving several GPUs through the C ABI (`stepth_multi`: a context and a host thread per device; what a Rust
    host links).  Pair batches are dealt to the devices without any collective; one giant pair runs in row bands."""

    def __init__(self, devices):
        self._lib = _ffi.lib()
        self.devices = [int(d) for d in devices]
        arr = (C.c_int32 * len(self.devices))(*self.devices)
        h = C.c_void_p()
        rc = self._lib.stepth_cuda_multi_create(arr, len(self.devices), C.byref(h))
        if rc != _ffi.OK:
            raise _ffi.StepthError(rc, (self._lib.stepth_cuda_multi_last_error(None) or b"").decode())
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.stepth_cuda_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def launch_count(self) -> int:
        return int(self._lib.stepth_cuda_multi_launch_count(self._h))

    def _check(self, rc: int):
        if rc == _ffi.OK:
            return
        msg = (self._lib.stepth_cuda_multi_last_error(self._h) or b"").decode()
        if rc in _PANIC_TEXT:
            raise StepthPanic(rc, _PANIC_TEXT[rc])
        if rc in (_ffi.ERR_INVALID_ARG, _ffi.ERR_SIZE_MISMATCH):
            raise StepthIOError(msg or "invalid argument")
        raise _ffi.StepthError(rc, msg)

    def depth_dense_batch(self, mains, adds, precision, max_rings=255, flags=_ffi.RESAMPLE_U8, out=None):
        """Context.depth_dense_batch over all the devices: pair i runs on device i mod n"""
        n = len(mains)
        if n == 0 or len(adds) != n:
            raise StepthIOError("Sizes don't match")
        h, w = mains[0].shape[:2]
        ha, wa = adds[0].shape[:2]
        for m, a in zip(mains, adds):
            if tuple(m.shape) != (h, w, 4) or tuple(a.shape) != (ha, wa, 3):
                raise StepthIOError("Sizes don't match")
        p = _ffi.make_params(precision, max_rings, flags)
        out = out or {}
        depth = out.get("depth") or [np.empty((h, w), np.uint8) for _ in range(n)]
        mask = (out.get("mask") or [np.empty((h, w), np.uint8) for _ in range(n)]) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = (out.get("rgba") or [np.empty((h, w, 4), np.uint8) for _ in range(n)]) if flags & _ffi.FLAG_APPLY else None
        ft = np.zeros((n, 2), np.uint8)
        status = np.zeros(n, np.int32)

        def parr(seq):
            return None if seq is None else (C.c_void_p * n)(*[_ffi.ptr(x) for x in seq])

        self._check(self._lib.stepth_cuda_multi_depth_dense_batch(self._h, n, parr(mains), w, h, parr(adds), wa, ha, C.byref(p),
                                                                  parr(depth), parr(rgba), parr(mask), _ffi.ptr(ft), _ffi.ptr(status)))
        return {"depth": depth, "mask": mask, "rgba": rgba, "from_to": ft, "status": status}

    def depth_dense_banded(self, main_rgba, add_rgb, precision, max_rings=255, flags=_ffi.RESAMPLE_U8, out=None):
        """Context.depth_dense for ONE pair in row bands over the devices (host buffers in and out)"""
        main_rgba, add_rgb = _u8c(main_rgba, (4,)), _u8c(add_rgb, (3,))
        h, w = main_rgba.shape[:2]
        ha, wa = add_rgb.shape[:2]
        p = _ffi.make_params(precision, max_rings, flags)
        out = out or {}
        res = {"depth": out.get("depth") if out.get("depth") is not None else np.empty((h, w), np.uint8)}
        mask = (out.get("mask") if out.get("mask") is not None else np.empty((h, w), np.uint8)) if flags & _ffi.FLAG_FOREGROUND else None
        rgba = (out.get("rgba") if out.get("rgba") is not None else np.empty((h, w, 4), np.uint8)) if flags & _ffi.FLAG_APPLY else None
        ft = np.zeros(2, np.uint8)
        self._check(self._lib.stepth_cuda_multi_depth_dense_banded(self._h, _ffi.ptr(main_rgba), w, h, _ffi.ptr(add_rgb), wa, ha,
                                                                   C.byref(p), _ffi.ptr(res["depth"]), _ffi.ptr(rgba), _ffi.ptr(mask),
                                                                   _ffi.ptr(ft)))
        if mask is not None:
            res["mask"], res["from_to"] = mask, (int(ft[0]), int(ft[1]))
        if rgba is not None:
            res["rgba"] = rgba
        return res


def _max_splits(w: int, h: int) -> int:
    """(W*H as f32).log2().ceil()   depth_image.rs:101-103"""
    return int(np.ceil(np.log2(np.float32(w * h))))


def discretise_v0(main_px, precision, min_splits: int = 16, max_splits: int | None = None) -> np.ndarray:
    """Host stand-in for disage::open::rgb_discrete (depth_image.rs:104-109); NOT disage (source absent)."""
    main_px = _u8c(main_px)
    h, w, stride = main_px.shape
    if max_splits is None:
        max_splits = _max_splits(w, h)
    prec = np.ascontiguousarray(np.asarray(precision, dtype=np.uint8).reshape(3))
    L = _ffi.lib()
    n = L.stepth_host_discretise_v0(_ffi.ptr(main_px), stride, w, h, _ffi.ptr(prec), min_splits, max_splits, None, 0)
    blocks = np.zeros(n, dtype=BLOCK_DTYPE)
    L.stepth_host_discretise_v0(_ffi.ptr(main_px), stride, w, h, _ffi.ptr(prec), min_splits, max_splits,
                                blocks.ctypes.data, n)
    return blocks


_default = None
_default_lock = threading.Lock()


def default_context() -> Context:
    global _default
    with _default_lock:
        if _default is None:
            _default = Context()
        return _default
