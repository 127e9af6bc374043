#!/usr/bin/env python
"""bench.py -- Mpx*disparities/s of the depth-from-additional hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--kind nat|geo|worst]

One "step" = one pass of the hot path (ring search -> normalise -> unit-scale Gaussian -> Luma8 depth [-> invert ->
select_foreground -> apply_mask], depth_image.rs:110-134,220-245, mask_image.rs:205-213) over one batch of synthetic
stereo pairs.  The DEFAULT workload is BASELINE.json configs[3], the configuration the north-star's scaling target is
stated on: 64 pairs of 3840x2160, 256 rings, fused tail, dealt to the N ranks ("scaling": "strong").  The line also
carries `other_workloads`: configs[0] (the README pair, block mode and dense), configs[1] (1080p / 128 rings),
configs[2] (4K fused, 4 pairs) and configs[4] (one 16384x8192 pair in row bands), each with its own ms_per_step, e2e,
roofline, cpu_baseline and parity.

  value      W*H*D * pairs * K / t, inputs resident in HBM, device-timed (CUDA events, max over ranks)
  e2e        the same metric through the reference-facing C ABI call with HOST buffers (pinned): host->device copies of
             every input and device->host copies of every result inside the timed region; `e2e.pageable` = the same
             call with ordinary (pageable) host memory, what a Rust Vec<u8> is
  roofline   the dominant kernel stage (the ring search): algorithmic bytes per launch / its mean launch duration,
             measured with CUDA events inside the timed region, against MEASURED_PEAKS.json's copy bandwidth; `issue` =
             the instruction-issue ceiling that actually binds it (from the committed ncu capture); `int` = SURVEY 8(d)'s
             integer half, probes counted on the device over the pairs that were timed
  parity     after the timed region: sampled queries + whole finished planes of the results that were just timed,
             against the CPU oracle (a port of the reference; upstream parity unpinned).  A
             mismatch makes the run exit non-zero.
  cpu_baseline  the CPU oracle timed on this box's cores on a bounded sample of the same workload (rank 0, N = 1 only)

`--impl reference` times that CPU port alone with the reference's own threading (len/8-sized chunks -> 8-9 rayon tasks,
depth_image.rs:111-112) on a bounded sample per step and prints the same JSON line with "impl": "reference".
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PRECISION = [36, 36, 36]           # u8::MAX / 7 (Readme.md:15)
METRIC = "Mpx*disparities/sec"
SUB = int(os.environ.get("STEPTH_BENCH_SUB", "32"))   # most pairs per launch sequence (bounds the scratch in HBM; 16 -> 32: 64 x 4K 36.6 -> 36.2 ms)

# name -> shape of the workload.  pairs: per GPU when scaling == "weak", in total when "strong"
WORKLOADS = {
    "batch4k": dict(w=3840, h=2160, d=256, pairs=64, fused=True, scaling="strong", mode="dense", data="synthetic"),   # configs[3]
    "1080p": dict(w=1920, h=1080, d=128, pairs=16, fused=False, scaling="weak", mode="dense", data="synthetic"),     # configs[1]
    "4k_fused": dict(w=3840, h=2160, d=256, pairs=4, fused=True, scaling="weak", mode="dense", data="synthetic"),    # configs[2]
    "giant": dict(w=16384, h=8192, d=256, pairs=1, fused=False, scaling="strong", mode="bands", data="synthetic"),   # configs[4]
    "readme_dense": dict(w=600, h=400, d=255, pairs=1, fused=False, scaling="weak", mode="dense", data="assets"),    # configs[0]
    "readme_blocks": dict(w=600, h=400, d=255, pairs=1, fused=False, scaling="weak", mode="blocks", data="assets"),  # configs[0]
}
DEFAULT_WORKLOAD = "batch4k"
OTHER_WORKLOADS = ["readme_blocks", "readme_dense", "1080p", "4k_fused", "giant"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default="nat", choices=["nat", "geo", "worst"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--pairs", type=int, default=0, help="pairs per step (default: the workload's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="skip the CUDA-graph replay of the sub-millisecond workloads")
    ap.add_argument("--no-others", action="store_true", help="only the named workload, no other_workloads")
    ap.add_argument("--extra-kinds", action="store_true", help="also time the geo and worst distributions (N=1)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def reach(w, h, d):
    """the part of the additional image a query can touch: origins are half-coordinates (depth_image.rs:114-117)"""
    return min(w, w // 2 + d), min(h, h // 2 + d)


def search_alg_bytes(w, h, d, pairs):
    rw, rh = reach(w, h, d)
    return pairs * (4 * w * h + 3 * rw * rh + w * h)   # main RGBA read + reachable additional RGB read + d8 write


def ncu_profile(workload, kind):
    """what the committed `ncu --set full` capture of this workload's search stage says (profiles/traffic.json):
    {"traffic": dram bytes read + written per launch sequence, "issue": {...}} or None"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            v = json.load(f).get(f"{workload}_{kind}")
        if isinstance(v, (int, float)):
            return {"traffic": int(v)}
        return v
    return None


def bind_to_gpu_numa(index: int):
    """pins this process to the CPUs next to its GPU (NVML's ideal affinity) so that pinned host buffers and the threads that
    fill them sit on the GPU's NUMA node; returns the number of CPUs in the set (None when NVML cannot say)"""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return len(os.sched_getaffinity(0))
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while the timed regions run"""

    def __init__(self, index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.period = period_s
        self._stop_evt = threading.Event()
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path, timed on this box's cores (never the product path)
# ---------------------------------------------------------------------------------------------------------------
def load_assets():
    """the reference's README pair as decoded pixels (tests/golden/assets.npz, made by tests/golden/make_golden.py)"""
    import numpy as np

    g = np.load(os.path.join(ROOT, "tests", "golden", "assets.npz"))
    main, add = g["main"], g["additional"]
    rgba = np.ascontiguousarray(np.concatenate([main, np.full(main.shape[:2] + (1,), 255, np.uint8)], axis=-1))
    return rgba, np.ascontiguousarray(add)


def host_pair(wl, kind, seed):
    """one pair of the workload as numpy arrays, generated on the CPU (the reference arm must not need a GPU)"""
    if wl["data"] == "assets":
        return load_assets()
    from stepth_b200 import synth

    m, a = synth.make_pair(kind, wl["w"], wl["h"], seed=seed)
    return m.numpy(), a.numpy()


def cpu_pass_rows(oracle, wl, m, a, y0, rows, threads, blocks=None):
    """the reference pass (depth_image.rs:96-136 [+ tail]) on the CPU over main rows [y0, y0+rows) of one pair; the whole
    pair when rows >= h.  Returns the rows processed.  threads = 0: the reference's own len/8-sized chunks (8-9 tasks)."""
    import numpy as np

    w, h, d = wl["w"], wl["h"], wl["d"]
    if blocks is not None:
        oracle.depth_from_blocks(a, blocks, PRECISION, w, h, d, resample=oracle.RESAMPLE_U8, threads=threads)
        return h
    if rows >= h:
        depth = oracle.depth_dense(m, a, PRECISION, d, resample=oracle.RESAMPLE_U8, threads=threads)
    else:
        idx = np.arange(y0 * w, (y0 + rows) * w, dtype=np.uint32)
        sub, _ = oracle.search_dense_subset(m, a, PRECISION, idx, d, threads=threads)
        try:
            norm, _ = oracle.normalise(sub.reshape(rows, w))
        except oracle.OracleError:      # a band whose distances are all 0: the whole image would still normalise
            norm = sub.reshape(rows, w)
        depth = oracle.gaussian_resample_unit(norm, oracle.RESAMPLE_U8)
    if wl["fused"]:
        depth = oracle.invert_depth(depth)
        try:
            mask, _ = oracle.select_foreground(depth)
            oracle.apply_mask(m[y0:y0 + depth.shape[0]], mask)
        except oracle.OracleError:
            pass
    return min(rows, h)


def cpu_calibrate(oracle, wl, m, a, threads, target_s):
    """rows of one pair the CPU port gets through in about target_s (a multiple of 8, at most the whole pair)"""
    h = wl["h"]
    probe = min(h, 16)
    t0 = time.perf_counter()
    cpu_pass_rows(oracle, wl, m, a, (h // 2) // 8 * 8, probe, threads)
    dt = max(time.perf_counter() - t0, 1e-4)
    rows = int(probe * target_s / dt) // 8 * 8
    return h if rows >= h else max(8, rows)


def cpu_baseline(wl, name, kind, pair=None, budget_s=12.0):
    """bounded sample of the workload on all host cores: whole pairs when one fits the budget, else a row band"""
    from oracle import oracle

    cores = os.cpu_count() or 1
    m, a = pair if pair is not None else host_pair(wl, kind, 1)
    w, h, d = wl["w"], wl["h"], wl["d"]
    if wl["mode"] == "blocks":
        import stepth_b200 as sb

        blocks = sb.discretise_v0(m, PRECISION)
        t0 = time.perf_counter()
        n = 0
        while n < 3 and time.perf_counter() - t0 < budget_s:
            cpu_pass_rows(oracle, wl, m, a, 0, h, cores, blocks=blocks)
            n += 1
        dt = time.perf_counter() - t0
        return {"value": round(n * w * h * d / dt / 1e6, 3), "unit": METRIC, "cores": cores, "kind": "port",
                "sample": f"{n} x the whole {w}x{h} pair, {len(blocks)} blocks (disc_v0 stand-in, not timed), D={d}, "
                          f"oracle depth_from_blocks with {cores} threads, {dt:.1f} s"}
    rows = cpu_calibrate(oracle, wl, m, a, cores, budget_s * 0.8)
    y0 = 0 if rows >= h else ((h - rows) // 2) // 8 * 8
    t0 = time.perf_counter()
    done = cpu_pass_rows(oracle, wl, m, a, y0, rows, cores)
    dt = time.perf_counter() - t0
    what = f"one whole {w}x{h} '{kind}' pair" if rows >= h else f"main rows [{y0}, {y0 + rows}) of one {w}x{h} '{kind}' pair"
    return {"value": round(done * w * d / dt / 1e6, 3), "unit": METRIC, "cores": cores, "kind": "port",
            "sample": f"{what}, D={d}, oracle with {cores} threads, {dt:.1f} s"}


def workload_config(name, wl, kind, world, pairs_total, pairs_this_gpu, l2_note, parallelism):
    return {"workload": f"{name}_d{wl['d']}_{kind if wl['data'] == 'synthetic' else 'assets'}", "w": wl["w"], "h": wl["h"],
            "max_rings": wl["d"], "precision": PRECISION, "pairs_per_step_total": pairs_total,
            "pairs_per_step_this_gpu": pairs_this_gpu, "mode": wl["mode"],
            "flags": "RESAMPLE_U8" + ("|INVERT|FOREGROUND|APPLY" if wl["fused"] else ""), "l2": l2_note,
            "parallelism": parallelism}


def describe(name, wl, kind, world):
    """the `config` object of a workload: the GPU arm and the reference arm print the same one"""
    pairs = wl["pairs"]
    w, h = wl["w"], wl["h"]
    if wl["mode"] == "bands":
        return workload_config(name, wl, kind, world, 1, 1, "inputs larger than L2",
                               f"row bands x{world}: allreduce(MAX) + 3-row halo exchange + all-gather of depth bands (NCCL, called by the library)")
    total = pairs if wl["scaling"] == "strong" else pairs * world
    mine = (pairs + world - 1) // world if wl["scaling"] == "strong" else pairs
    in_mb = mine * 7 * w * h / 1e6
    l2 = (f"inputs {in_mb:.0f} MB per step per GPU > 126 MB L2 (no flush needed)" if in_mb > 126
          else "L2 flushed between timed iterations (256 MB memset, outside the per-iteration events)")
    return workload_config(name, wl, kind, world, total, mine, l2, f"pair-batch x{world} (no collective)")


def run_reference(args):
    """--impl reference: the CPU port with the reference's own parallelism (8-9 chunks).  Each step is a bounded sample of
    the GPU arm's workload: main rows of one pair sized so that a step takes about 2 s (the whole pair when that fits)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle

    name, wl = args.workload, WORKLOADS[args.workload]
    w, h, d = wl["w"], wl["h"], wl["d"]
    cores = os.cpu_count() or 1
    n_pairs = 1 if wl["data"] == "assets" or wl["mode"] == "bands" else min(4, max(1, args.steps))
    pairs = [host_pair(wl, args.kind, i + 1) for i in range(n_pairs)]
    blocks = None
    if wl["mode"] == "blocks":
        import stepth_b200 as sb

        blocks = sb.discretise_v0(pairs[0][0], PRECISION)
        rows = h
    else:
        rows = cpu_calibrate(oracle, wl, *pairs[0], 0, 1.5)
    n_bands = max(1, h // rows)

    def one(i):
        m, a = pairs[i % n_pairs]
        y0 = ((i // n_pairs) * 5 % n_bands) * rows if rows < h else 0     # bands spread over the image
        return cpu_pass_rows(oracle, wl, m, a, y0, rows, 0, blocks=blocks)  # threads = 0 -> len/8-sized chunks = 8-9 threads

    for i in range(args.warmup):
        one(i)
    t0 = time.perf_counter()
    done = 0
    for i in range(args.steps):
        done += one(args.warmup + i)
    dt = time.perf_counter() - t0
    val = done * w * d / dt / 1e6
    used = min(9, cores)
    sample = (f"{args.steps} steps x " + ("one whole pair" if rows >= h else f"{rows} main rows of one {w}x{h} pair (bands spread over the image, {n_pairs} pairs)")
              + f", reference-faithful 8-9 chunk threads on {cores} cores")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": METRIC, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "u8", "data": wl["data"],
        "config": describe(name, wl, args.kind, max(1, args.gpus)),
        "cpu_baseline": {"value": round(val, 3), "unit": METRIC, "cores": used, "kind": "port", "sample": sample},
        "e2e": {"value": round(val, 3), "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------------------
# GPU arm: timing
# ---------------------------------------------------------------------------------------------------------------
class Env:
    """what every workload runner needs: torch, the process group, this rank's context"""

    def __init__(self, args):
        import torch

        self.args, self.torch = args, torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- stepth_b200 has no CPU fallback (use --impl reference for the CPU port)")
        self.numa_cpus = bind_to_gpu_numa(self.local)   # before any pinned allocation
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        import stepth_b200 as sb

        self.sb = sb
        self.ctx = sb.Context(self.local)
        self.dev = torch.device("cuda", self.local)
        self.flush_buf = None

    def flush_l2(self):
        """writes a buffer twice the size of L2 (126 MB) so that nothing of the previous iteration is cache-resident"""
        if self.flush_buf is None:
            self.flush_buf = self.torch.empty(256 << 20, dtype=self.torch.uint8, device=self.dev)
        self.flush_buf.zero_()

    def max_over_ranks(self, x: float) -> float:
        if self.world > 1:
            t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def sum_over_ranks(self, x: float) -> float:
        if self.world > 1:
            t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
            return float(t.item())
        return x

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()


def time_resident(env, step, steps, warmup, flush=False):
    """K timed steps over HBM-resident inputs; `step(events)` enqueues one step and records (start, end) CUDA events around
    its ring-search stage.  Returns (total ms [max over ranks], search-stage ms summed over the steps).  flush: the inputs
    are smaller than L2 -> every iteration is timed by its own event pair with an L2 flush in between."""
    torch = env.torch
    for _ in range(warmup):
        step(None)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    env.barrier()
    if flush:
        it = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for i in range(steps):
            env.flush_l2()
            it[i][0].record()
            step(ev[i])
            it[i][1].record()
        torch.cuda.synchronize()
        total = sum(a.elapsed_time(b) for a, b in it)
    else:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(steps):
            step(ev[i])
        t1.record()
        torch.cuda.synchronize()
        total = t0.elapsed_time(t1)
    env.barrier()
    search = sum(a.elapsed_time(b) for a, b in ev)
    return env.max_over_ranks(total), search


def time_graph_replay(env, step, steps, results, want):
    """A sub-millisecond resident step is a dozen short launches with gaps between them.  The stage API only enqueues work on the
    caller's stream and, after its first call, neither allocates nor synchronises, so a caller can capture the step ONCE into a
    CUDA graph and replay it (what a serving loop would do).  `results()` -> the tensors the step writes, `want` = their values
    after the eager run.  Returns {"ms_per_step", "bit_exact_vs_eager", ..} -- or {"error": ..}: this never fails the run."""
    torch = env.torch
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step(None)                       # the library's scratch for this shape exists before the capture starts
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
            step(None)
        torch.cuda.synchronize()
        for t in results():
            t.zero_()                        # the replay has to produce the results, not inherit them
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        same = all(bool(torch.equal(a, b)) for a, b in zip(results(), want))
        it = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for i in range(steps):
            env.flush_l2()
            it[i][0].record()
            g.replay()
            it[i][1].record()
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in it) / steps
        del g
        return {"ms_per_step": round(ms, 4), "steps": steps, "bit_exact_vs_eager": same,
                "note": "the same resident step captured once into a CUDA graph by the caller and replayed (L2 flushed between replays)"}
    except Exception as e:   # noqa: BLE001 -- a measurement aid: the line is printed without it
        try:
            torch.cuda.synchronize()
        except Exception:   # noqa: BLE001
            pass
        return {"error": f"{type(e).__name__}: {str(e)[:200]}"}


def time_e2e(env, step, steps, warmup, flush=False):
    """`step()` = the C-ABI host-buffer call: H2D of all inputs + the whole path + D2H of the results (synchronous);
    wall clock around the calls, max over ranks"""
    for _ in range(max(1, min(warmup, 2))):
        step()
    env.barrier()
    if flush:
        dt = 0.0
        for _ in range(steps):
            env.flush_l2()
            env.torch.cuda.synchronize()
            t0 = time.perf_counter()
            step()
            dt += (time.perf_counter() - t0) * 1e3
    else:
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        dt = (time.perf_counter() - t0) * 1e3
    return env.max_over_ranks(dt)


def link_ceiling(env, h2d_bytes, d2h_bytes):
    """the end-to-end roofline: ms this rank's share of one step's host<->device traffic takes with NO kernels -- page-locked
    memory, uploads and downloads on two streams at once, every rank at the same time (max over ranks, best of 3)"""
    torch = env.torch
    piece = 32 << 20
    cap = 768 << 20
    h_in = torch.empty(max(1, min(h2d_bytes, cap)), dtype=torch.uint8).pin_memory()
    h_out = torch.empty(max(1, min(d2h_bytes, cap)), dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty_like(h_in, device=env.dev), torch.empty_like(h_out, device=env.dev)
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

    def run():
        env.barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(s_up):
            for o in range(0, h2d_bytes, piece):
                a = o % h_in.numel()
                n = min(piece, h_in.numel() - a, h2d_bytes - o)
                d_in[a:a + n].copy_(h_in[a:a + n], non_blocking=True)
        with torch.cuda.stream(s_down):
            for o in range(0, d2h_bytes, piece):
                a = o % h_out.numel()
                n = min(piece, h_out.numel() - a, d2h_bytes - o)
                h_out[a:a + n].copy_(d_out[a:a + n], non_blocking=True)
        torch.cuda.synchronize()
        return env.max_over_ranks((time.perf_counter() - t0) * 1e3)

    run()
    return min(run() for _ in range(3))


def build_roofline(name, kind, wl, alg_bytes, pairs_per_launch, search_launch_ms, share, kernel_desc):
    hbm, which = peaks()
    achieved = alg_bytes / (search_launch_ms * 1e-3) / 1e9
    prof = ncu_profile(name, kind) or {}
    traffic = prof.get("traffic")
    if traffic is not None and prof.get("pairs_per_launch"):   # the capture's launch sequence may hold another number of pairs
        traffic = int(traffic * pairs_per_launch / prof["pairs_per_launch"])
    r = {"bound": "hbm", "achieved": round(achieved, 2), "peak": hbm, "unit": "GB/s", "frac": round(achieved / hbm, 5),
         "traffic": traffic, "peak_source": which, "kernel": kernel_desc,
         "alg_bytes_per_launch": alg_bytes, "pairs_per_launch": pairs_per_launch, "launch_ms": round(search_launch_ms, 4),
         "share_of_step": round(share, 4),
         "note": "the stage is instruction-issue / latency bound, not HBM bound (DESIGN.md section 5): `issue` is the ceiling "
                 "that binds it; frac is reported as the contract asks"}
    if prof.get("issue"):
        r["issue"] = prof["issue"]   # {"kernel", "frac", "lanes_active", "inst_executed", "duration_us", "source"} from profiles/
    return r


def int_roofline(env, wl, mains, adds, n_first, search_launch_ms, alg_bytes):
    """SURVEY 8(d)'s other half: O_alg = 2 integer ops per probe of the reference order, the probes counted ON THE DEVICE
    (in-order kernel with a counter) over the very pairs of the timed launch sequence; P_INT measured by a register-only kernel"""
    from stepth_b200 import device

    torch = env.torch
    hbm, _ = peaks()
    _, probes = device.count_probes(env.ctx, mains[:n_first], adds[:n_first], PRECISION, wl["d"])
    torch.cuda.synchronize()
    per_q = (probes.double() / float(wl["w"] * wl["h"])).cpu().tolist()
    total = float(probes.sum().item())
    p_mixed, p_logic = env.ctx.int_peak(False), env.ctx.int_peak(True)
    ops = 2.0 * total
    t_int_ms, t_hbm_ms = ops / p_mixed * 1e3, alg_bytes / (hbm * 1e9) * 1e3
    return {
        "bound": "int-alu", "unit": "Tiop/s", "ops_per_launch": round(ops),
        "probes_per_query_min": round(min(per_q), 2), "probes_per_query_mean": round(sum(per_q) / len(per_q), 2),
        "probes_per_query_max": round(max(per_q), 2), "pairs_counted": n_first,
        # early exit makes the nominal W*H*D much larger than what the reference order really visits
        "reference_probes_per_s": round(total / (search_launch_ms * 1e-3), 1),
        "nominal_over_effective": round(n_first * wl["w"] * wl["h"] * (2.0 * wl["d"] - 1) ** 2 / max(total, 1.0), 2),
        "achieved": round(ops / (search_launch_ms * 1e-3) / 1e12, 3), "peak": round(p_mixed / 1e12, 3),
        "frac": round(ops / (search_launch_ms * 1e-3) / p_mixed, 4), "peak_logic_only": round(p_logic / 1e12, 3),
        "peak_source": "measured (stepth_cuda_int_peak: LOP3+IMAD.IADD on both integer pipes; logic_only = LOP3+SHF, alu pipe)",
        "t_roof_ms": round(max(t_int_ms, t_hbm_ms), 4), "t_hbm_ms": round(t_hbm_ms, 4), "t_int_ms": round(t_int_ms, 4),
        "frac_of_t_roof": round(max(t_int_ms, t_hbm_ms) / search_launch_ms, 4),
        "note": "O_alg = 2 ops x the probes the reference order needs, counted by stepth_cuda_dev_count_probes over every pair of "
                "the timed launch sequence; the tile search skips probes, so this is an equivalent-work figure, not a "
                "utilisation, and may exceed 1 (SURVEY.md 8(d))"}


# ---------------------------------------------------------------------------------------------------------------
# GPU arm: parity of what was just timed, against the CPU oracle (test infrastructure; the checker, never the path)
# ---------------------------------------------------------------------------------------------------------------
ORACLE_NOTE = "port of the reference (oracle/stepth_oracle.c), upstream parity unpinned"


def parity_dense(env, wl, m_t, a_t, raw_t, depth_t, mx, mask_t=None, rgba_t=None, from_to=None, host_depth=None,
                 n_queries=2000, seed=0):
    """one pair: sampled raw distances vs the oracle's search; the WHOLE finished plane (+ mask, + rgba) vs the oracle's
    normalise / Gaussian / invert / select_foreground / apply_mask of those raw distances; the host-buffer call's plane too"""
    import numpy as np

    from oracle import oracle

    w, h, d = wl["w"], wl["h"], wl["d"]
    m, a, raw = m_t.cpu().numpy(), a_t.cpu().numpy(), raw_t.cpu().numpy()
    rng = np.random.default_rng(1234 + seed)
    idx = rng.choice(w * h, min(n_queries, w * h), replace=False).astype(np.uint32)
    sub, _ = oracle.search_dense_subset(m, a, PRECISION, idx, d, threads=os.cpu_count() or 1)
    ok = bool(np.array_equal(raw.reshape(-1)[idx], sub))
    planes = 0
    norm, mx_ref = oracle.normalise(raw)
    ok = ok and mx_ref == int(mx)
    depth = oracle.gaussian_resample_unit(norm, oracle.RESAMPLE_U8)
    if wl["fused"]:
        depth = oracle.invert_depth(depth)
    ok = ok and bool(np.array_equal(depth_t.cpu().numpy(), depth))
    planes += 1
    if host_depth is not None:
        ok = ok and bool(np.array_equal(np.asarray(host_depth), depth))
        planes += 1
    if wl["fused"]:
        mask, ft = oracle.select_foreground(depth)
        ok = ok and tuple(int(v) for v in from_to) == tuple(ft)
        ok = ok and bool(np.array_equal(mask_t.cpu().numpy(), mask))
        ok = ok and bool(np.array_equal(rgba_t.cpu().numpy(), oracle.apply_mask(m, mask)))
        planes += 2
    return {"bit_exact": ok, "checked_queries": int(len(idx)), "checked_planes": planes}


def finish_parity(env, p):
    """all ranks agree on the verdict; counts are summed"""
    ok = env.sum_over_ranks(0.0 if p["bit_exact"] else 1.0) == 0.0
    return {"bit_exact": ok, "checked_queries": int(env.sum_over_ranks(p["checked_queries"])),
            "checked_planes": int(env.sum_over_ranks(p["checked_planes"])), "oracle": ORACLE_NOTE}


# ---------------------------------------------------------------------------------------------------------------
# GPU arm: dense pair batches (configs[1], [2], [3] and the README pair in dense mode)
# ---------------------------------------------------------------------------------------------------------------
def run_dense(env, name, kind, steps, warmup, main_line):
    """-> the JSON line of one dense workload (rank 0 uses it; every rank must call it)"""
    import numpy as np

    from stepth_b200 import bands, device, synth

    torch, args, sb, ctx, dev = env.torch, env.args, env.sb, env.ctx, env.dev
    wl = dict(WORKLOADS[name])
    if main_line and args.pairs:
        wl["pairs"] = args.pairs
    W, H, D = wl["w"], wl["h"], wl["d"]
    rank, world = env.rank, env.world
    flags = sb.RESAMPLE_U8 | ((sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY) if wl["fused"] else 0)
    # pairs this rank processes per step: weak = a fixed batch per GPU; strong = the workload's pairs dealt to the ranks
    if wl["scaling"] == "strong":
        my_pairs = bands.shard_pairs(wl["pairs"], rank, world)
        total_pairs = wl["pairs"]
    else:
        my_pairs = [rank * wl["pairs"] + i for i in range(wl["pairs"])]
        total_pairs = wl["pairs"] * world
    B = len(my_pairs)
    sub = min(B, SUB)
    in_mb = B * 7 * W * H / 1e6
    flush = in_mb <= 126
    if flush:
        steps = max(steps, 20)   # a sub-millisecond step: more iterations, each with its own event pair and an L2 flush before it

    def make_batch(k):
        if wl["data"] == "assets":
            m, a = load_assets()
            return torch.from_numpy(m).to(dev)[None].contiguous(), torch.from_numpy(a).to(dev)[None].contiguous()
        ms, as_ = [], []
        for i in my_pairs:
            m, a = synth.make_pair(k, W, H, seed=i + 1, device=dev)
            ms.append(m)
            as_.append(a)
        return torch.stack(ms).contiguous(), torch.stack(as_).contiguous()

    mains, adds = make_batch(kind)
    pipes = {}

    def pipe_for(n):
        if n not in pipes:
            pipes[n] = device.DensePipeline(ctx, n, W, H, W, H, PRECISION, D, flags)
        return pipes[n]

    def make_step(mains, adds):
        def step(events):
            for j, i0 in enumerate(range(0, B, sub)):
                n = min(sub, B - i0)
                # the search-stage events bracket the first launch sequence of the step (all are alike)
                pipe_for(n).run(mains[i0:i0 + n], adds[i0:i0 + n], search_events=events if j == 0 else None)
        return step

    step = make_step(mains, adds)
    l0 = ctx.launch_count()
    step(None)
    torch.cuda.synchronize()
    launches_per_step = ctx.launch_count() - l0
    for p in pipes.values():
        assert int(p.status.abs().sum()) == 0, "a pair of the bench batch failed (max == 0?)"

    sampler = ClockSampler(env.local) if main_line else None
    if sampler:
        sampler.start()
    total_ms, search_ms = time_resident(env, step, steps, warmup, flush=flush)
    clocks = sampler.stop() if sampler else None
    px_disp = W * H * D
    value = total_pairs * steps * px_disp / (total_ms * 1e-3) / 1e6
    search_launch_ms = search_ms / steps
    alg = search_alg_bytes(W, H, D, sub)
    n_seq = (B + sub - 1) // sub
    roofline = build_roofline(name, kind, wl, alg, sub, search_launch_ms, search_ms * n_seq / total_ms,
                              "ring search stage (stepth_cuda_dev_search_dense: tile_planes + coarse_transpose + search_near + search_far)")

    # the results of the LAST pair of this rank's batch, as the resident pipeline left them (checked below)
    last = B - 1
    lp = pipe_for(B - (n_seq - 1) * sub)
    lj = last - (n_seq - 1) * sub
    keep = {"raw": lp.d8[lj].clone(), "depth": lp.depth[lj].clone(), "mx": int(lp.max[lj])}
    if wl["fused"]:
        keep.update(mask=lp.mask[lj].clone(), rgba=lp.rgba_out[lj].clone(), ft=lp.ranges[lj, :2].tolist())

    e2e, host_last = None, None
    if not args.no_e2e:
        host_m = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(B)]
        host_a = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(B)]
        out = {"depth": [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)]}
        if wl["fused"]:
            out["mask"] = [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)]
            out["rgba"] = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(B)]
        for i in range(B):
            host_m[i].copy_(mains[i])
            host_a[i].copy_(adds[i])

        def e2e_step(hm=host_m, ha=host_a, ho=out, n=B):
            for i0 in range(0, n, sub):
                sl = slice(i0, min(n, i0 + sub))
                ctx.depth_dense_batch(hm[sl], ha[sl], PRECISION, D, flags, out={k: v[sl] for k, v in ho.items()})

        e2e_steps = max(2, min(steps, 5)) if not flush else max(steps, 10)
        e2e_ms = time_e2e(env, e2e_step, e2e_steps, warmup, flush=flush)
        rw, rh = reach(W, H, D)
        h2d, d2h = B * (4 * W * H + 3 * rw * rh), B * W * H * (6 if wl["fused"] else 1)
        e2e = {"value": round(total_pairs * e2e_steps * px_disp / (e2e_ms * 1e-3) / 1e6, 3), "unit": METRIC,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": round(e2e_ms / e2e_steps, 3), "steps": e2e_steps,
               "link_gbs_this_gpu": round((h2d + d2h) / (e2e_ms / e2e_steps * 1e-3) / 1e9, 2),
               "api": "stepth_cuda_depth_dense_batch (host buffers, pinned; chunks pipelined over copy/compute/copy streams)"}
        if main_line:
            floor_ms = link_ceiling(env, h2d, d2h)
            tot = env.sum_over_ranks(h2d + d2h)
            e2e["link_floor_ms"] = round(floor_ms, 3)
            e2e["link_ceiling_gbs"] = round(tot / (floor_ms * 1e-3) / 1e9, 2)        # all ranks together, no kernels
            e2e["frac_of_link_ceiling"] = round(floor_ms / (e2e_ms / e2e_steps), 4)
        host_last = out["depth"][last].numpy().copy()
        # the same call with ordinary pageable memory (what a Rust Vec<u8> is): the library stages it through its own
        # pinned ring.  Bounded to one launch sequence of pairs.
        nb = min(B, sub)
        pg_m = [np.array(host_m[i].numpy(), copy=True) for i in range(nb)]
        pg_a = [np.array(host_a[i].numpy(), copy=True) for i in range(nb)]
        pg_out = {k: [np.empty(tuple(v[0].shape), np.uint8) for _ in range(nb)] for k, v in out.items()}
        pg_steps = max(2, min(steps, 4)) if not flush else max(steps, 10)
        pg_ms = time_e2e(env, lambda: e2e_step(pg_m, pg_a, pg_out, nb), pg_steps, warmup, flush=flush)
        pg_total = nb * world if wl["scaling"] == "weak" or world == 1 else int(env.sum_over_ranks(nb))
        e2e["pageable"] = {"value": round(pg_total * pg_steps * px_disp / (pg_ms * 1e-3) / 1e6, 3),
                           "ms_per_step": round(pg_ms / pg_steps, 3), "pairs_per_step_this_gpu": nb, "steps": pg_steps,
                           "note": "numpy (malloc) buffers in and out; staged through the library's pinned ring by its copy threads"}
        e2e["pageable"]["vs_pinned"] = round(e2e["pageable"]["value"] / e2e["value"], 3)   # both are pairs/s x px_disp
        same = all(np.array_equal(pg_out["depth"][i], out["depth"][i].numpy()) for i in range(nb))
        assert same, "pageable and pinned host-buffer calls disagree"
        del host_m, host_a, out, pg_m, pg_a, pg_out

    parity = None
    if not args.no_parity:
        p = parity_dense(env, wl, mains[last], adds[last], keep["raw"], keep["depth"], keep["mx"], keep.get("mask"),
                         keep.get("rgba"), keep.get("ft"), host_depth=host_last, seed=rank)
        parity = finish_parity(env, p)
        parity["what"] = f"pair {my_pairs[last]} (the last of each rank's batch): sampled raw distances + whole finished planes, resident and host-buffer results"

    graph = None
    if flush and world == 1 and n_seq == 1 and not args.no_graph:
        want = [keep["raw"], keep["depth"]] + ([keep["mask"], keep["rgba"]] if wl["fused"] else [])
        graph = time_graph_replay(env, step, steps, lambda: [lp.d8[lj], lp.depth[lj]] + ([lp.mask[lj], lp.rgba_out[lj]] if wl["fused"] else []), want)

    extra = {}
    if main_line and args.extra_kinds and world == 1:
        for k2 in ("geo", "worst"):
            if k2 == kind:
                continue
            m2, a2 = make_batch(k2)
            steps2 = max(1, min(steps, 3))
            t_ms, s_ms = time_resident(env, make_step(m2, a2), steps2, 1)
            extra[k2] = {"value": round(total_pairs * steps2 * px_disp / (t_ms * 1e-3) / 1e6, 3), "ms_per_step": round(t_ms / steps2, 3),
                         "search_ms_per_launch": round(s_ms / steps2, 3)}
            del m2, a2

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, name, kind, pair=(mains[0].cpu().numpy(), adds[0].cpu().numpy()))
        if wl["data"] == "synthetic":
            roofline["int"] = int_roofline(env, wl, mains, adds, sub, search_launch_ms, alg)

    line = {
        "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": round(total_ms / steps, 4), "higher_is_better": True,
        "scaling": wl["scaling"], "vs_baseline": None, "dtype": "u8", "data": wl["data"],
        "config": describe(name, wl, kind, world),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "parity": parity, "gpu_launches": launches_per_step * steps,
    }
    if graph is not None:
        line["graph_replay"] = graph
    if clocks is not None:
        line["clocks"] = clocks
    if extra:
        line["other_distributions"] = extra
    del mains, adds, pipes, keep
    torch.cuda.empty_cache()
    return line


# ---------------------------------------------------------------------------------------------------------------
# GPU arm: the README pair in block mode (configs[0]: the reference's actual call shape, depth_image.rs:104-122)
# ---------------------------------------------------------------------------------------------------------------
def run_blocks(env, name, kind, steps, warmup, main_line):
    import numpy as np

    from stepth_b200 import device

    torch, args, sb, ctx, dev = env.torch, env.args, env.sb, env.ctx, env.dev
    wl = WORKLOADS[name]
    W, H, D = wl["w"], wl["h"], wl["d"]
    m, a = load_assets()
    # disage's source is absent; disc_v0 is the documented stand-in for rgb_discrete (depth_image.rs:104-109).  It runs on the
    # GPU (csrc/disc.cuh) INSIDE the timed step: main image -> block list -> search -> paint -> finish.  The host version of the
    # same stand-in gives the list the CPU port is timed on (not timed there: the reference's discretiser is not ours to time).
    blocks = ctx.discretise_v0(m, PRECISION)
    nb = len(blocks)
    blocks_host = sb.discretise_v0(m, PRECISION)
    disc_equal = blocks.tobytes() == blocks_host.tobytes()
    blocks_t = torch.zeros((nb, 20), dtype=torch.uint8, device=dev)
    main_t = torch.from_numpy(m).to(dev).contiguous()
    add_t = torch.from_numpy(a).to(dev).contiguous()
    pipe = device.BlockPipeline(ctx, nb, W, H, a.shape[1], a.shape[0], PRECISION, D, sb.RESAMPLE_U8)

    def step(events):
        pipe.run(blocks_t, add_t, search_events=events, main_px=main_t)

    l0 = ctx.launch_count()
    step(None)
    torch.cuda.synchronize()
    launches_per_step = ctx.launch_count() - l0
    assert int(pipe.status[0]) == 0
    steps_r = max(steps, 20)
    total_ms, search_ms = time_resident(env, step, steps_r, warmup, flush=True)
    px_disp = W * H * D
    value = env.world * steps_r * px_disp / (total_ms * 1e-3) / 1e6
    # algorithmic bytes of the block search: 20 B per block read + the reachable additional image + 1 B per block written
    rw, rh = min(a.shape[1], (W + 1) // 2 + D + 1), min(a.shape[0], (H + 1) // 2 + D + 1)
    alg = nb * 21 + 3 * rw * rh
    roofline = build_roofline(name, "assets", wl, alg, 1, search_ms / steps_r, search_ms / total_ms,
                              "ring search stage, block mode (stepth_cuda_dev_search_blocks: tile_planes + coarse_transpose + search_near<blocks> + search_far<blocks>)")
    raw_gpu, depth_gpu = pipe.d8.cpu().numpy(), pipe.depth[0].cpu().numpy()
    disc_equal = disc_equal and int(pipe.n_blocks[0]) == nb and blocks_t.cpu().numpy().tobytes() == blocks_host.tobytes()

    e2e, host_depth = None, None
    if not args.no_e2e:
        pin_a = torch.from_numpy(a).pin_memory()
        pin_m = torch.from_numpy(m).pin_memory()
        res = {}

        def call(main_host, add_host):
            res["out"] = ctx.depth_from_additional(main_host, add_host, PRECISION, D, sb.RESAMPLE_U8)

        e2e_steps = max(steps, 20)
        ms = time_e2e(env, lambda: call(pin_m.numpy(), pin_a.numpy()), e2e_steps, warmup, flush=True)
        host_depth = res["out"]["depth"].copy()
        disc_equal = disc_equal and res["out"]["n_blocks"] == nb
        ms_pg = time_e2e(env, lambda: call(m, a), e2e_steps, warmup, flush=True)
        e2e = {"value": round(env.world * e2e_steps * px_disp / (ms * 1e-3) / 1e6, 3), "unit": METRIC,
               "h2d_bytes_per_step": 4 * W * H + 3 * rw * rh, "d2h_bytes_per_step": W * H + 4,
               "ms_per_step": round(ms / e2e_steps, 4), "steps": e2e_steps,
               "api": "stepth_cuda_depth_from_additional (host buffers: main RGBA + additional RGB in, pinned; discretiser, search, "
                      "paint and finish on the device; the depth plane lands in a numpy array)",
               "pageable": {"value": round(env.world * e2e_steps * px_disp / (ms_pg * 1e-3) / 1e6, 3),
                            "ms_per_step": round(ms_pg / e2e_steps, 4), "steps": e2e_steps}}

    parity = None
    if not args.no_parity:
        from oracle import oracle

        ref_depth, ref_raw = oracle.depth_from_blocks(a, blocks, PRECISION, W, H, D, resample=oracle.RESAMPLE_U8,
                                                      threads=os.cpu_count() or 1, want_raw=True)
        ok = bool(np.array_equal(raw_gpu, ref_raw)) and bool(np.array_equal(depth_gpu, ref_depth)) and disc_equal
        planes = 1
        if host_depth is not None:
            ok = ok and bool(np.array_equal(host_depth, ref_depth))
            planes += 1
        parity = finish_parity(env, {"bit_exact": ok, "checked_queries": nb, "checked_planes": planes})
        parity["what"] = ("the device discretiser's block list against the host stand-in's (bytes), every block's raw distance + the "
                          "whole finished plane, resident and host-buffer results")
        parity["block_list_identical"] = bool(disc_equal)

    graph = None
    if env.world == 1 and not args.no_graph:
        want = [pipe.d8.clone(), pipe.depth.clone(), blocks_t.clone()]
        graph = time_graph_replay(env, step, steps_r, lambda: [pipe.d8, pipe.depth, blocks_t], want)

    cpu = None
    if env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, name, kind, pair=(m, a), budget_s=4.0)
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": env.world, "steps": steps_r, "warmup": warmup,
        "ms_per_step": round(total_ms / steps_r, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "assets", "config": dict(describe(name, wl, kind, env.world), n_blocks=nb,
                                                      discretiser="disc_v0 stand-in (disage source absent) on the device, inside the timed step and the e2e call"),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "parity": parity, "gpu_launches": launches_per_step * steps_r,
    }
    if graph is not None:
        line["graph_replay"] = graph
    del pipe, blocks_t, add_t, main_t
    torch.cuda.empty_cache()
    return line


# ---------------------------------------------------------------------------------------------------------------
# GPU arm: one giant pair in row bands (configs[4])
# ---------------------------------------------------------------------------------------------------------------
def run_giant(env, name, kind, steps, warmup, main_line):
    """ONE pair, main rows banded over the ranks.  The orchestration lives below the C ABI (csrc/multi_api.cuh): this rank's
    stepth_comm + stepth_cuda_dev_depth_dense_band = ring search of the band, MAX all-reduce of the normaliser, 3-row halo
    exchange, finish, and an all-gather of the depth bands so that every GPU ends with the assembled map (NCCL, called by the
    library); bands.CommBandRunner only allocates the tensors."""
    import numpy as np

    from stepth_b200 import bands, device, synth

    torch, args, sb, ctx, dev, dist = env.torch, env.args, env.sb, env.ctx, env.dev, env.dist
    wl = WORKLOADS[name]
    W, H, D = wl["w"], wl["h"], wl["d"]
    rank, world = env.rank, env.world
    flags = sb.RESAMPLE_U8
    runner = bands.CommBandRunner(ctx, W, H, H, PRECISION, D, flags)
    b = runner.band
    # every rank generates the same pair from the same seed and keeps only its rows (synthetic data; no scatter timed)
    m, a = synth.make_pair(kind, W, H, seed=1, device=dev)
    mb, ab = m[None, b.y0:b.y1].contiguous(), a[None, b.ay0:b.ay1].contiguous()
    want_cpu_copy = not args.no_parity or (rank == 0 and world == 1 and not args.no_cpu_baseline)
    m_np = m.cpu().numpy() if want_cpu_copy else None
    a_np = a.cpu().numpy() if want_cpu_copy else None
    del m, a
    torch.cuda.empty_cache()
    gather = world > 1
    l0 = ctx.launch_count()
    out = runner.run(mb, ab, gather_depth=gather)
    launches = ctx.launch_count() - l0
    assert out["status"] == 0

    def step(events):
        if events is not None:
            events[0].record()
        runner.enqueue(mb, ab, gather_depth=gather)
        if events is not None:
            events[1].record()

    sampler = ClockSampler(env.local) if main_line else None
    if sampler:
        sampler.start()
    total_ms, _ = time_resident(env, step, steps, warmup)
    clocks = sampler.stop() if sampler else None
    out = runner.result()
    assert out["status"] == 0
    value = steps * W * H * D / (total_ms * 1e-3) / 1e6
    # where each rank's time went in the last step (CUDA events recorded by the library between the stages)
    mine = runner.stage_ms()
    stage_ms = [mine]
    if world > 1:
        stage_ms = [None] * world
        dist.all_gather_object(stage_ms, mine)
    # roofline of the search stage on the slowest band: this rank's main rows + the additional rows it can reach
    rw = min(W, W // 2 + D)
    my_alg = 4 * W * (b.y1 - b.y0) + 3 * rw * (b.ay1 - b.ay0) + W * (b.y1 - b.y0)
    slow = env.max_over_ranks(mine.get("search", 0.0))
    alg_all = int(env.sum_over_ranks(my_alg))
    roofline = build_roofline(name, kind, wl, alg_all // world, 1, slow, slow / (total_ms / steps),
                              "ring search stage of one band (tile_planes + coarse_transpose + search_near + search_far), the slowest rank's time, mean band bytes")

    e2e = None
    if not args.no_e2e:
        hm, ha = mb.cpu().pin_memory(), ab.cpu().pin_memory()
        hd = torch.empty((1, b.y1 - b.y0, W), dtype=torch.uint8).pin_memory()
        dm, da = torch.empty_like(mb), torch.empty_like(ab)

        def e2e_step():
            dm.copy_(hm, non_blocking=True)
            da.copy_(ha, non_blocking=True)
            runner.enqueue(dm, da, gather_depth=False)      # bands go straight back to pinned host memory (no gather)
            hd.copy_(runner.depth_band, non_blocking=True)
            runner.result()                                 # waits for the stream

        e2e_steps = max(2, min(steps, 5))
        ms = time_e2e(env, e2e_step, e2e_steps, warmup)
        e2e = {"value": round(e2e_steps * W * H * D / (ms * 1e-3) / 1e6, 3), "unit": METRIC,
               "h2d_bytes_per_step": int(env.sum_over_ranks(hm.numel() + ha.numel())),
               "d2h_bytes_per_step": int(env.sum_over_ranks(hd.numel())),
               "ms_per_step": round(ms / e2e_steps, 3), "steps": e2e_steps,
               "api": "stepth_cuda_dev_depth_dense_band per rank: its band (+ reachable additional rows) up from pinned memory, "
                      "MAX all-reduce + halo exchange inside the library, its depth band back to pinned memory"}
        del hm, ha, hd, dm, da

    parity = None
    if not args.no_parity:
        from oracle import oracle

        out = runner.run(mb, ab, gather_depth=gather)
        rows = b.y1 - b.y0
        # the raw distances of this band, by the very search call the band function makes
        raw_t, _ = device.search_dense(ctx, mb, ab, PRECISION, D, 0, y0=b.y0, add_y0=b.ay0, add_img_h=H, img_h=H)
        raw_band = raw_t[0].cpu().numpy()
        rng = np.random.default_rng(77 + rank)
        n_q = min(2000, rows * W)
        local = rng.choice(rows * W, n_q, replace=False).astype(np.int64)
        local[: min(200, n_q)] %= min(rows, 3) * W                      # some right under the upper seam
        idx = (local + b.y0 * W).astype(np.uint32)
        sub, _ = oracle.search_dense_subset(m_np, a_np, PRECISION, idx, D, threads=max(1, (os.cpu_count() or 1) // world))
        ok = bool(np.array_equal(raw_band.reshape(-1)[local], sub))
        planes = 0
        # the finished plane, whole: rank 0 assembles the raw bands and runs the oracle's normalise + Gaussian over them
        depth_t = out["depth_band"].contiguous()
        if world > 1:
            assert len({p.y1 - p.y0 for p in runner.plan}) == 1, "equal bands expected for the parity gather"
            raws = torch.empty((world,) + tuple(raw_t.shape), dtype=torch.uint8, device=dev)
            deps = torch.empty_like(raws)
            dist.all_gather_into_tensor(raws, raw_t.contiguous())
            dist.all_gather_into_tensor(deps, depth_t)
            ok = ok and bool(torch.equal(deps.reshape(H, W), out["depth"].reshape(H, W)))   # the library's own all-gather, on every rank
            raw_full, depth_full = raws.reshape(H, W), deps.reshape(H, W)
        else:
            raw_full, depth_full = raw_t.reshape(H, W), depth_t.reshape(H, W)
        if rank == 0:
            norm, _ = oracle.normalise(raw_full.cpu().numpy())
            ok = ok and bool(np.array_equal(depth_full.cpu().numpy(), oracle.gaussian_resample_unit(norm, oracle.RESAMPLE_U8)))
            planes = 1
        parity = finish_parity(env, {"bit_exact": ok, "checked_queries": n_q, "checked_planes": planes})
        parity["what"] = "every band: sampled raw distances (some at the seams) vs the oracle on the whole pair; the assembled finished plane, whole"

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, name, kind, pair=(m_np, a_np), budget_s=8.0)
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": round(total_ms / steps, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(describe(name, wl, kind, world), bands=[[p.y0, p.y1, p.ay0, p.ay1] for p in runner.plan],
                       exchange={1: "peer windows: the library's own kernels store maxima, halo rows and the depth bands into the other "
                                    "ranks' HBM over NVLink (csrc/peer.cuh); NCCL only bootstraps the window handles",
                                 0: "NCCL (all-reduce MAX, send/recv halos, all-gather of the depth bands)"}.get(
                                     runner.peer_mode() if world > 1 and runner.peer_mode() >= 0 else (0 if world > 1 else -1), "one rank: none")),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "parity": parity, "gpu_launches": launches * steps,
        "stage_ms_per_rank": stage_ms,
    }
    if clocks is not None:
        line["clocks"] = clocks
    runner.close()
    del mb, ab, out, runner
    torch.cuda.empty_cache()
    return line


RUNNERS = {"dense": run_dense, "blocks": run_blocks, "bands": run_giant}


def compact(line):
    """an other_workloads entry: the same keys as a line, minus what only the main line needs"""
    return {k: line[k] for k in ("value", "ms_per_step", "steps", "scaling", "config", "roofline", "cpu_baseline", "e2e", "parity",
                                 "gpu_launches", "stage_ms_per_rank", "graph_replay") if k in line}


def run_ours(args):
    env = Env(args)
    wl = WORKLOADS[args.workload]
    line = RUNNERS[wl["mode"]](env, args.workload, args.kind, args.steps, args.warmup, True)
    line["host"] = {"cpus": os.cpu_count(), "cpus_bound_to_this_gpu": env.numa_cpus}
    bad = [args.workload] if line.get("parity") and not line["parity"]["bit_exact"] else []
    if not args.no_others:
        others = {}
        for name in OTHER_WORKLOADS:
            if name == args.workload:
                continue
            sub_line = RUNNERS[WORKLOADS[name]["mode"]](env, name, args.kind, max(3, min(args.steps, 10)), 3, False)
            others[name] = compact(sub_line)
            if sub_line.get("parity") and not sub_line["parity"]["bit_exact"]:
                bad.append(name)
        line["other_workloads"] = others
    if env.rank == 0:
        emit(line)
    if env.world > 1:
        env.dist.destroy_process_group()
    if bad:
        sys.stderr.write(f"bench.py: PARITY FAILURE against the oracle in {bad}\n")
        raise SystemExit(3)


_STDOUT_FD = None


def quiet_stdout():
    """stdout carries ONE JSON line: libraries that printf there (NCCL's version banner at communicator creation) are sent to
    stderr for the length of the run; `emit` writes the line to the real stdout."""
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    text = (json.dumps(line) + "\n").encode()
    os.write(_STDOUT_FD if _STDOUT_FD is not None else 1, text)


def main():
    args = parse_args()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
