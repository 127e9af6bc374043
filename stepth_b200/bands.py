"""Multi-GPU orchestration of the hot path: one process per GPU (`torch.distributed`), SURVEY.md 8(e).

Two ways the path shards:

* **pair batches** (BASELINE.json configs[3]): pairs are independent -> `shard_pairs` deals pair i to rank i % world;
  no data-path collective at all.
* **row bands of one giant pair** (configs[4]): rank r owns main rows [y0, y1).  Origins are half-coordinates
  (depth_image.rs:114-117), so it needs additional rows [(y0+1)/2 - max, y1/2 + max] -- `band_plan`.  The path has
  three real exchange steps, all tiny except the last:
    1. all-reduce MAX of the normaliser (depth_image.rs:124): 1 u32 per pair;
    2. the unit-scale Gaussian (:133) reads 3 rows across each seam: neighbours exchange 3-row halos of the raw
       distance plane (all-gather of 6 edge rows per rank);
    3. select_foreground (:220-223): all-reduce SUM of the 256-bin histogram; every rank then runs the same k-means;
    4. optional all-gather of the depth bands when the assembled map must live on every GPU (NCCL over NVLink);
       otherwise each band goes straight back to its rank's pinned host memory.

`BandRunner` holds the exchange logic and is compute-agnostic: the product passes `DeviceStages` (the CUDA stage API,
stepth_b200.device); the CPU test-suite passes a stand-in with the same four methods and runs the very same exchange
code over `gloo` with world_size 2 (tests/test_bands_gloo.py).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

from . import _ffi

HALO = 3  # rows of the Gaussian's support on each side (image 0.23.14: support 3.0 at unit scale)


def shard_pairs(n_pairs: int, rank: int, world: int) -> list[int]:
    """pair i -> rank i % world (config 4); every rank gets floor or ceil of n/world pairs"""
    return list(range(rank, n_pairs, world))


@dataclass(frozen=True)
class Band:
    rank: int
    y0: int      # first main row of the band
    y1: int      # one past the last main row
    ay0: int     # first additional row the band can probe
    ay1: int     # one past the last


def band_plan(h: int, add_h: int, max_rings: int, world: int) -> list[Band]:
    """Cuts main rows [0, h) into `world` contiguous bands (multiples of 2 rows so that the two main rows that share an
    origin row stay together) and gives each the additional rows it can reach: origin rows (y+1)/2 for y in [y0, y1)
    -> [(y0+1)/2, y1/2], widened by max_rings - 1 on both sides.  ONE implementation: the library's
    `stepth_host_band_plan` (csrc/multi.cuh), which the C++ band path uses too."""
    if world < 1 or h < 1:
        raise ValueError("need at least one rank and one row")
    import ctypes as C

    lib = _ffi.lib()
    bands = []
    out = (C.c_uint32 * 4)()
    for r in range(world):
        lib.stepth_host_band_plan(h, add_h, max_rings, world, r, C.byref(out))
        bands.append(Band(r, int(out[0]), int(out[1]), int(out[2]), int(out[3])))
    return bands


class DeviceStages:
    """The CUDA stage API (stepth_cuda_dev_*) behind the four calls BandRunner makes."""

    def __init__(self, ctx):
        self.ctx = ctx

    def search(self, main_band, add_band, precision, max_rings, flags, y0, add_y0, add_img_h, img_h):
        from . import device

        return device.search_dense(self.ctx, main_band, add_band, precision, max_rings, flags & _ffi.FLAG_PIXELOPPS_SAT,
                                   y0=y0, add_y0=add_y0, add_img_h=add_img_h, img_h=img_h)

    def finish(self, src, mx, img_h, flags, src_y0, out_y0, out_rows, hist):
        from . import device

        return device.finish(self.ctx, src, mx, img_h, flags, src_y0=src_y0, out_y0=out_y0, out_rows=out_rows, hist=hist)

    def depth_split(self, hist):
        from . import device

        return device.depth_split(self.ctx, hist, 2)

    def mask_apply(self, depth, from_to, rgba):
        from . import device

        return device.mask_apply(self.ctx, depth, from_to, rgba, want_mask=True)


class BandRunner:
    """depth_image.rs:110-134 (+ fused tail) for ONE pair whose main rows are banded over the ranks of `group`."""

    def __init__(self, stages, w: int, h: int, add_h: int, precision, max_rings: int, flags: int, group=None):
        self.st = stages
        self.w, self.h, self.add_h = int(w), int(h), int(add_h)
        self.precision, self.max_rings, self.flags = list(precision), int(max_rings), int(flags)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.plan = band_plan(self.h, self.add_h, self.max_rings, self.world)
        self.band = self.plan[self.rank]
        self.profile = False      # True: `run` brackets its stages with CUDA events and leaves the times in `stage_ms`
        self.stage_ms: dict[str, float] = {}
        self._marks: list = []

    def _mark(self, name: str) -> None:
        if self.profile:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._marks.append((name, ev))

    def _close_marks(self) -> None:
        if self.profile and self._marks:
            torch.cuda.synchronize()
            self.stage_ms = {b[0]: round(a[1].elapsed_time(b[1]), 4) for a, b in zip(self._marks, self._marks[1:])}
            self._marks = []

    # ---- exchange steps ------------------------------------------------------------------------------------------
    def _allreduce(self, t: torch.Tensor, op) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=op, group=self.group)
        return t

    def exchange_halos(self, d8_band: torch.Tensor) -> tuple[torch.Tensor, int]:
        """-> (raw distance rows [src_y0, src_y0 + rows) = own band plus up to 3 rows of each neighbour, src_y0)."""
        b = self.band
        n, rows, w = d8_band.shape
        if self.world == 1:
            return d8_band, b.y0
        # every rank contributes its first and last HALO rows (zero-padded when the band is shorter)
        edge = torch.zeros((2, n, HALO, w), dtype=torch.uint8, device=d8_band.device)
        k = min(HALO, rows)
        if k:
            edge[0, :, :k] = d8_band[:, :k]
            edge[1, :, HALO - k:] = d8_band[:, rows - k:]
        edges = [torch.empty_like(edge) for _ in range(self.world)]
        dist.all_gather(edges, edge, group=self.group)
        # rows above: walk up through the previous ranks' LAST rows until HALO rows are collected (bands may be short)
        above, below = [], []
        need = min(HALO, b.y0)
        r = self.rank - 1
        while need > 0 and r >= 0:
            have = min(HALO, self.plan[r].y1 - self.plan[r].y0, need)
            if have:
                above.insert(0, edges[r][1][:, HALO - have:])
                need -= have
            r -= 1
        need = min(HALO, self.h - b.y1)
        r = self.rank + 1
        while need > 0 and r < self.world:
            have = min(HALO, self.plan[r].y1 - self.plan[r].y0, need)
            if have:
                below.append(edges[r][0][:, :have])
                need -= have
            r += 1
        parts = above + [d8_band] + below
        src = torch.cat(parts, dim=1).contiguous() if len(parts) > 1 else d8_band
        src_y0 = b.y0 - sum(p.shape[1] for p in above)
        return src, src_y0

    # ---- the path ---------------------------------------------------------------------------------------------------
    def run(self, main_band: torch.Tensor, add_band: torch.Tensor, gather_depth: bool = False) -> dict:
        """main_band [1, y1-y0, w, 4], add_band [1, ay1-ay0, wa, 3] of THIS rank (see `band_plan`).
        Returns depth_band [1, rows, w] (+ mask_band, rgba_band, from_to when the tail is fused; + depth [1, h, w]
        on every rank when gather_depth)."""
        b, f = self.band, self.flags
        dev = main_band.device
        rows = b.y1 - b.y0
        self._mark("start")
        if rows > 0:
            d8, mx = self.st.search(main_band, add_band, self.precision, self.max_rings, f, b.y0, b.ay0, self.add_h, self.h)
            mx = mx.to(torch.int32)
        else:
            d8 = torch.empty((1, 0, self.w), dtype=torch.uint8, device=dev)
            mx = torch.zeros(1, dtype=torch.int32, device=dev)
        self._mark("search")
        mx = self._allreduce(mx, dist.ReduceOp.MAX)                      # exchange 1: the normaliser
        src, src_y0 = self.exchange_halos(d8)                            # exchange 2: 3-row halos
        self._mark("allreduce_max+halos")
        want_fg = bool(f & _ffi.FLAG_FOREGROUND)
        hist = torch.zeros((1, 256), dtype=torch.int32, device=dev) if want_fg else None
        out = {}
        if rows > 0:
            depth, status = self.st.finish(src, mx, self.h, f & (_ffi.RESAMPLE_U8 | _ffi.RESAMPLE_F32 | _ffi.FLAG_INVERT),
                                           src_y0, b.y0, rows, hist)
        else:
            depth = torch.empty((1, 0, self.w), dtype=torch.uint8, device=dev)
            status = torch.zeros(1, dtype=torch.int32, device=dev)
        # max == 0 is a property of the whole image: every rank sees the same all-reduced value.  Without the fused tail
        # nothing on the host depends on the status, so it is read once, after the last exchange has been enqueued (a
        # blocking read here would expose every later launch and collective to the host's latency).
        def read_status() -> int:
            return _ffi.ERR_DIV_BY_ZERO_MAX if int(mx[0]) == 0 else int(status[0]) if rows > 0 else 0

        out["depth_band"] = depth
        out["d8_band"] = d8          # raw distances of this band (parity checks read them; nothing else does)
        self._mark("finish")
        if want_fg:
            out["status"] = read_status()
        if want_fg and out["status"] == 0:
            hist = self._allreduce(hist, dist.ReduceOp.SUM)              # exchange 3: the histogram
            ranges, n_out, st2 = self.st.depth_split(hist)
            out["status"] = int(st2[0])
            if out["status"] == 0:
                out["from_to"] = (int(ranges[0, 0]), int(ranges[0, 1]))
                if rows > 0:
                    rgba = main_band if f & _ffi.FLAG_APPLY else None
                    mask, rgba_out = self.st.mask_apply(depth, ranges, rgba)
                    out["mask_band"], out["rgba_band"] = mask, rgba_out
        if gather_depth:                                                 # exchange 4: assemble the map everywhere
            out["depth"] = self.gather_bands(depth)
        if not want_fg:
            out["status"] = read_status()
        self._mark("tail+gather")
        self._close_marks()
        return out

    def gather_bands(self, band: torch.Tensor) -> torch.Tensor:
        """all-gather of (possibly unequal) row bands -> [n, h, w] on every rank"""
        if self.world == 1:
            return band
        n, _, w = band.shape
        rows = [p.y1 - p.y0 for p in self.plan]
        if n == 1 and len(set(rows)) == 1:
            # equal bands of one pair are consecutive slabs of the assembled map: gather straight into it
            out = torch.empty((1, self.h, w), dtype=band.dtype, device=band.device)
            dist.all_gather_into_tensor(out.view(-1), band.contiguous().view(-1), group=self.group)
            return out
        max_rows = max(rows)
        padded = torch.zeros((n, max_rows, w), dtype=band.dtype, device=band.device)
        padded[:, : band.shape[1]] = band
        parts = [torch.empty_like(padded) for _ in range(self.world)]
        dist.all_gather(parts, padded, group=self.group)
        return torch.cat([parts[p.rank][:, : p.y1 - p.y0] for p in self.plan], dim=1).contiguous()


class _DeviceBytes:
    """a device buffer owned by the library, shown to torch through __cuda_array_interface__ (no copy)"""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


class CommBandRunner:
    """The row-band path below the C ABI (csrc/multi_api.cuh): this rank's `stepth_comm` (context + NCCL communicator created
    from a 128-byte id that rank 0 makes and torch.distributed broadcasts) and `stepth_cuda_dev_depth_dense_band`, which
    enqueues the search, the MAX all-reduce, the halo exchange, finish, the histogram all-reduce + tail and the band
    all-gather on torch's current stream.  Python only allocates the tensors."""

    def __init__(self, ctx, w: int, h: int, add_h: int, precision, max_rings: int, flags: int, wa: int | None = None, group=None):
        import ctypes as C

        self.ctx, self.C = ctx, C
        self.w, self.h, self.add_h, self.wa = int(w), int(h), int(add_h), int(wa or w)
        self.flags = int(flags)
        self.params = _ffi.make_params(precision, max_rings, flags)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.plan = band_plan(self.h, self.add_h, int(max_rings), self.world)
        self.band = self.plan[self.rank]
        lib = ctx._lib
        uid = torch.zeros(128, dtype=torch.uint8)
        if self.world > 1:
            if self.rank == 0:
                buf = (C.c_uint8 * 128)()
                rc = lib.stepth_cuda_comm_unique_id(C.addressof(buf))
                if rc:
                    raise _ffi.StepthError(rc, lib.stepth_cuda_last_error(None).decode())
                uid = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
            dev = torch.device("cuda", ctx.device)
            t = uid.to(dev)
            dist.broadcast(t, src=0, group=group)
            uid = t.cpu()
        self._uid = uid.numpy().copy()
        handle = C.c_void_p()
        ctx.check(lib.stepth_cuda_comm_create(ctx.handle, self.world, self.rank, self._uid.ctypes.data, C.byref(handle)))
        self._h = handle
        dev = torch.device("cuda", ctx.device)
        rows = self.band.y1 - self.band.y0
        u8 = dict(dtype=torch.uint8, device=dev)
        self.depth_band = torch.empty((1, rows, self.w), **u8)
        self.fg, self.apply = bool(flags & _ffi.FLAG_FOREGROUND), bool(flags & _ffi.FLAG_APPLY)
        self.mask_band = torch.empty((1, rows, self.w), **u8) if self.fg else None
        self.rgba_band = torch.empty((1, rows, self.w, 4), **u8) if self.apply else None
        self.depth_full = None

    def close(self):
        if getattr(self, "_h", None):
            self.ctx._lib.stepth_cuda_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def enqueue(self, main_band: torch.Tensor, add_band: torch.Tensor, gather_depth: bool = False) -> None:
        """main_band [1, rows, w, 4], add_band [1, arows, wa, 3] of THIS rank (device tensors); asynchronous"""
        b = self.band
        assert tuple(main_band.shape) == (1, b.y1 - b.y0, self.w, 4) and tuple(add_band.shape) == (1, b.ay1 - b.ay0, self.wa, 3)
        st = torch.cuda.current_stream().cuda_stream
        if gather_depth and self.depth_full is None:
            # the assembled map lives in this rank's peer window when the group has one (csrc/peer.cuh: the other ranks' finish
            # kernels store their bands straight into it over NVLink); a tensor of our own when the exchanges go through NCCL
            ptr = self.C.c_void_p()
            self.ctx.check(self.ctx._lib.stepth_cuda_comm_full_map(self._h, st, self.w, self.h, self.add_h, int(self.params.max_rings),
                                                                   self.C.byref(ptr)))
            if ptr.value:
                self._window_full = _DeviceBytes(ptr.value, (1, self.h, self.w))
                self.depth_full = torch.as_tensor(self._window_full, device=main_band.device)
                assert self.depth_full.data_ptr() == ptr.value
            else:
                self.depth_full = torch.empty((1, self.h, self.w), dtype=torch.uint8, device=main_band.device)
        self.ctx.check(self.ctx._lib.stepth_cuda_dev_depth_dense_band(
            self._h, st, self.w, self.h, self.wa, self.add_h, self.C.byref(self.params), _ffi.ptr(main_band), _ffi.ptr(add_band),
            _ffi.ptr(self.depth_band), _ffi.ptr(self.depth_full) if gather_depth else None, _ffi.ptr(self.mask_band),
            _ffi.ptr(self.rgba_band)))

    def result(self) -> dict:
        """waits for the stream; -> status (+ from_to when the tail is fused)"""
        C = self.C
        status, ft = C.c_int32(0), (C.c_uint8 * 2)()
        st = torch.cuda.current_stream().cuda_stream
        self.ctx.check(self.ctx._lib.stepth_cuda_comm_band_result(self._h, st, C.byref(status), C.addressof(ft)))
        out = {"status": int(status.value), "depth_band": self.depth_band}
        if self.fg and out["status"] == 0:
            out.update(from_to=(int(ft[0]), int(ft[1])), mask_band=self.mask_band, rgba_band=self.rgba_band)
        if self.depth_full is not None:
            out["depth"] = self.depth_full
        return out

    def run(self, main_band, add_band, gather_depth: bool = False) -> dict:
        self.enqueue(main_band, add_band, gather_depth)
        return self.result()

    def peer_mode(self) -> int:
        """1: the exchanges are stores into the peers' HBM over NVLink (csrc/peer.cuh); 0: NCCL; -1: not decided yet"""
        return int(self.ctx._lib.stepth_cuda_comm_peer_mode(self._h))

    def stage_ms(self) -> dict:
        """device time of the last call per stage (after result())"""
        ms = (self.C.c_float * 4)()
        self.ctx.check(self.ctx._lib.stepth_cuda_comm_band_stage_ms(self._h, self.C.byref(ms)))
        return {k: round(float(v), 4) for k, v in zip(("search", "allreduce_max+halos", "finish", "tail+gather"), ms)}
