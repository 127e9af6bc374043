"""Device-resident driver of the stage API (`stepth_cuda_dev_*`): inputs and outputs are torch CUDA tensors,
work is enqueued on torch's current stream, nothing is copied to the host.  torch is plumbing here (device
memory + streams); every kernel that runs is one of libstepth_cuda.so's.

Used by bench.py (the HBM-resident `value`), by the multi-GPU orchestration and by the GPU parity tests.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _ffi
from .context import Context


def _stream_handle() -> int:
    return torch.cuda.current_stream().cuda_stream


class DensePipeline:
    """depth_image.rs:110-134 (+ fused invert / select_foreground / apply_mask) for a batch of same-shape pairs
    that already sit in HBM: main_rgba [n,h,w,4] u8, add_rgb [n,ha,wa,3] u8."""

    def __init__(self, ctx: Context, n_pairs: int, w: int, h: int, wa: int, ha: int, precision, max_rings: int = 255,
                 flags: int = _ffi.RESAMPLE_U8, want_mask: bool = True):
        self.ctx = ctx
        self.n, self.w, self.h, self.wa, self.ha = int(n_pairs), int(w), int(h), int(wa), int(ha)
        self.flags = int(flags)
        self.params = _ffi.make_params(precision, max_rings, flags)
        dev = torch.device("cuda", ctx.device)
        u8 = dict(dtype=torch.uint8, device=dev)
        self.d8 = torch.empty((self.n, h, w), **u8)
        self.depth = torch.empty((self.n, h, w), **u8)
        self.fg = bool(flags & _ffi.FLAG_FOREGROUND)
        self.apply = bool(flags & _ffi.FLAG_APPLY)
        self.mask = torch.empty((self.n, h, w), **u8) if self.fg and want_mask else None
        self.rgba_out = torch.empty((self.n, h, w, 4), **u8) if self.apply else None
        # small per-pair state
        self.max = torch.zeros(self.n, dtype=torch.int32, device=dev)
        self.hist = torch.zeros((self.n, 256), dtype=torch.int32, device=dev)
        self.status = torch.zeros(self.n, dtype=torch.int32, device=dev)
        self.n_out = torch.zeros(self.n, dtype=torch.int32, device=dev)
        self.ranges = torch.zeros((self.n, 512), **u8)
        self.shape = _ffi.DevShape(self.n, self.w, self.h, self.h, 0, self.wa, self.ha, 0, self.ha)

    def run(self, main_rgba: torch.Tensor, add_rgb: torch.Tensor, search_events=None) -> None:
        """Enqueue the whole path on the current stream (asynchronous).  `search_events` = (start, end) CUDA events
        recorded around the ring-search stage (bench.py's roofline timing)."""
        assert main_rgba.is_cuda and add_rgb.is_cuda and main_rgba.dtype == torch.uint8 and add_rgb.dtype == torch.uint8
        assert tuple(main_rgba.shape) == (self.n, self.h, self.w, 4) and tuple(add_rgb.shape) == (self.n, self.ha, self.wa, 3)
        lib, ctx, st = self.ctx._lib, self.ctx, _stream_handle()
        p = C.byref(self.params)
        if self.fg:
            self.hist.zero_()
        self.status.zero_()
        if search_events is not None:
            search_events[0].record()
        ctx.check(lib.stepth_cuda_dev_search_dense(ctx.handle, st, C.byref(self.shape), p, _ffi.ptr(main_rgba),
                                                   _ffi.ptr(add_rgb), _ffi.ptr(self.d8), _ffi.ptr(self.max)))
        if search_events is not None:
            search_events[1].record()
        ctx.check(lib.stepth_cuda_dev_finish(ctx.handle, st, self.n, self.w, self.h, _ffi.ptr(self.d8), 0, self.h,
                                             _ffi.ptr(self.max), self.flags, _ffi.ptr(self.depth), 0, self.h,
                                             _ffi.ptr(self.hist) if self.fg else None, _ffi.ptr(self.status)))
        if self.fg:
            ctx.check(lib.stepth_cuda_dev_depth_split(ctx.handle, st, self.n, _ffi.ptr(self.hist), 2, _ffi.ptr(self.ranges),
                                                      _ffi.ptr(self.n_out), _ffi.ptr(self.status)))
            if self.mask is not None or self.apply:
                ctx.check(lib.stepth_cuda_dev_mask_apply(ctx.handle, st, self.n, self.w * self.h, _ffi.ptr(self.depth),
                                                         _ffi.ptr(self.ranges), 512, _ffi.ptr(self.status),
                                                         _ffi.ptr(main_rgba) if self.apply else None,
                                                         _ffi.ptr(self.mask), _ffi.ptr(self.rgba_out)))


class BlockPipeline:
    """depth_image.rs:110-134 from a DiscretePixel list that already sits in HBM (block mode, one pair):
    blocks [n_blocks] (20-byte records, as a u8 tensor [n_blocks, 20]), add_rgb [ha, wa, 3]."""

    def __init__(self, ctx: Context, n_blocks: int, w: int, h: int, wa: int, ha: int, precision, max_rings: int = 255,
                 flags: int = _ffi.RESAMPLE_U8):
        self.ctx, self.nb, self.w, self.h, self.wa, self.ha = ctx, int(n_blocks), int(w), int(h), int(wa), int(ha)
        self.flags = int(flags)
        self.params = _ffi.make_params(precision, max_rings, flags)
        dev = torch.device("cuda", ctx.device)
        u8 = dict(dtype=torch.uint8, device=dev)
        self.d8 = torch.empty(self.nb, **u8)
        self.plane = torch.empty((1, h, w), **u8)
        self.depth = torch.empty((1, h, w), **u8)
        self.max = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)

    def discretise(self, main_px: torch.Tensor, blocks: torch.Tensor, min_splits: int = 16, max_splits: int | None = None):
        """the stand-in discretiser on the device (csrc/disc.cuh): main_px [h, w, 3|4] -> blocks [nb, 20]; the count lands in
        self.n_blocks (a device u32; the caller sized `blocks` from an earlier count)"""
        assert main_px.is_cuda and main_px.dtype == torch.uint8 and tuple(main_px.shape[:2]) == (self.h, self.w)
        if max_splits is None:
            from .context import _max_splits

            max_splits = _max_splits(self.w, self.h)
        if not hasattr(self, "n_blocks"):
            self.n_blocks = torch.zeros(1, dtype=torch.int32, device=main_px.device)
        self.ctx.check(self.ctx._lib.stepth_cuda_dev_discretise_v0(
            self.ctx.handle, _stream_handle(), _ffi.ptr(main_px), int(main_px.shape[2]), self.w, self.h, self.params.precision,
            min_splits, max_splits, _ffi.ptr(blocks), int(blocks.shape[0]), _ffi.ptr(self.n_blocks)))

    def run(self, blocks: torch.Tensor, add_rgb: torch.Tensor, search_events=None, main_px: torch.Tensor | None = None) -> None:
        """main_px given: the block list is rebuilt from it on the device first (the whole of load_depth_from_additional)"""
        assert blocks.is_cuda and blocks.dtype == torch.uint8 and tuple(blocks.shape) == (self.nb, 20)
        assert tuple(add_rgb.shape) == (self.ha, self.wa, 3)
        lib, ctx, st = self.ctx._lib, self.ctx, _stream_handle()
        p = C.byref(self.params)
        self.status.zero_()
        if main_px is not None:
            self.discretise(main_px, blocks)
        if search_events is not None:
            search_events[0].record()
        ctx.check(lib.stepth_cuda_dev_search_blocks(ctx.handle, st, p, _ffi.ptr(add_rgb), self.wa, self.ha, _ffi.ptr(blocks),
                                                    self.nb, _ffi.ptr(self.d8), _ffi.ptr(self.max)))
        if search_events is not None:
            search_events[1].record()
        ctx.check(lib.stepth_cuda_dev_paint_blocks(ctx.handle, st, _ffi.ptr(blocks), _ffi.ptr(self.d8), self.nb, self.w, self.h,
                                                   _ffi.ptr(self.plane)))
        ctx.check(lib.stepth_cuda_dev_finish(ctx.handle, st, 1, self.w, self.h, _ffi.ptr(self.plane), 0, self.h,
                                             _ffi.ptr(self.max), self.flags, _ffi.ptr(self.depth), 0, self.h, None,
                                             _ffi.ptr(self.status)))


def search_dense(ctx: Context, main_rgba: torch.Tensor, add_rgb: torch.Tensor, precision, max_rings=255, flags=0,
                 y0: int = 0, add_y0: int = 0, add_img_h: int | None = None, img_h: int | None = None):
    """K1 only: -> (d8 [n,h,w] u8, max [n] i32).  Band geometry (y0 / add_y0 / add_img_h) is for row-banded runs."""
    n, h, w, _ = main_rgba.shape
    _, ha, wa, _ = add_rgb.shape
    d8 = torch.empty((n, h, w), dtype=torch.uint8, device=main_rgba.device)
    mx = torch.zeros(n, dtype=torch.int32, device=main_rgba.device)
    shape = _ffi.DevShape(n, w, h, img_h or h, y0, wa, ha, add_y0, add_img_h if add_img_h is not None else ha)
    p = _ffi.make_params(precision, max_rings, flags)
    ctx.check(ctx._lib.stepth_cuda_dev_search_dense(ctx.handle, _stream_handle(), C.byref(shape), C.byref(p),
                                                    _ffi.ptr(main_rgba), _ffi.ptr(add_rgb), _ffi.ptr(d8), _ffi.ptr(mx)))
    return d8, mx


def finish(ctx: Context, src: torch.Tensor, mx: torch.Tensor, img_h: int, flags: int, src_y0: int = 0, out_y0: int = 0,
           out_rows: int | None = None, hist: torch.Tensor | None = None):
    """K2 on a band: src [n, src_rows, w] holds global rows [src_y0, ..); -> (depth [n,out_rows,w], status [n])."""
    n, src_rows, w = src.shape
    out_rows = src_rows if out_rows is None else out_rows
    depth = torch.empty((n, out_rows, w), dtype=torch.uint8, device=src.device)
    status = torch.zeros(n, dtype=torch.int32, device=src.device)
    ctx.check(ctx._lib.stepth_cuda_dev_finish(ctx.handle, _stream_handle(), n, w, img_h, _ffi.ptr(src), src_y0, src_rows,
                                              _ffi.ptr(mx), flags, _ffi.ptr(depth), out_y0, out_rows, _ffi.ptr(hist),
                                              _ffi.ptr(status)))
    return depth, status


def depth_split(ctx: Context, hist: torch.Tensor, zones: int = 2):
    n = hist.shape[0]
    ranges = torch.zeros((n, 512), dtype=torch.uint8, device=hist.device)
    n_out = torch.zeros(n, dtype=torch.int32, device=hist.device)
    status = torch.zeros(n, dtype=torch.int32, device=hist.device)
    ctx.check(ctx._lib.stepth_cuda_dev_depth_split(ctx.handle, _stream_handle(), n, _ffi.ptr(hist), zones, _ffi.ptr(ranges),
                                                   _ffi.ptr(n_out), _ffi.ptr(status)))
    return ranges, n_out, status


def mask_apply(ctx: Context, depth: torch.Tensor, from_to: torch.Tensor, rgba: torch.Tensor | None = None,
               want_mask: bool = True):
    """K4: depth [n, ...] u8, from_to [n, >=2] u8 -> (mask | None, rgba_out | None)"""
    n = depth.shape[0]
    px = depth[0].numel()
    mask = torch.empty_like(depth) if want_mask else None
    out = torch.empty_like(rgba) if rgba is not None else None
    ctx.check(ctx._lib.stepth_cuda_dev_mask_apply(ctx.handle, _stream_handle(), n, px, _ffi.ptr(depth), _ffi.ptr(from_to),
                                                  from_to.stride(0), None, _ffi.ptr(rgba), _ffi.ptr(mask), _ffi.ptr(out)))
    return mask, out


def histogram(ctx: Context, data: torch.Tensor) -> torch.Tensor:
    hist = torch.zeros(256, dtype=torch.int32, device=data.device)
    ctx.check(ctx._lib.stepth_cuda_dev_histogram(ctx.handle, _stream_handle(), _ffi.ptr(data), data.numel(), _ffi.ptr(hist)))
    return hist


def count_probes(ctx: Context, main_rgba: torch.Tensor, add_rgb: torch.Tensor, precision, max_rings=255, flags=0):
    """Measurement aid: the in-order kernel with a probe counter -> (d8 [n,h,w], probes [n] i64): per pair, the in-bounds
    pixels the reference's scan order (helpers.rs:26-47) visits up to and including each query's first match."""
    n, h, w, _ = main_rgba.shape
    _, ha, wa, _ = add_rgb.shape
    d8 = torch.empty((n, h, w), dtype=torch.uint8, device=main_rgba.device)
    mx = torch.zeros(n, dtype=torch.int32, device=main_rgba.device)
    probes = torch.zeros(n, dtype=torch.int64, device=main_rgba.device)
    shape = _ffi.DevShape(n, w, h, h, 0, wa, ha, 0, ha)
    p = _ffi.make_params(precision, max_rings, flags)
    ctx.check(ctx._lib.stepth_cuda_dev_count_probes(ctx.handle, _stream_handle(), C.byref(shape), C.byref(p), _ffi.ptr(main_rgba),
                                                    _ffi.ptr(add_rgb), _ffi.ptr(d8), _ffi.ptr(mx), _ffi.ptr(probes)))
    return d8, probes
