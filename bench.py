#!/usr/bin/env python
"""bench.py -- Mpx*disparities/s of the depth-from-additional hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--kind nat|geo|worst] [--pairs B]

One "step" = one pass of the hot path (ring search -> normalise -> unit-scale Gaussian -> Luma8 depth,
depth_image.rs:110-134) over one batch of B synthetic stereo pairs of BASELINE.json's configs[1] shape
(1920x1080 RGB pair, max_rings = 128, precision [36;3]); the batch is larger than L2 (B = 16 -> 232 MB of inputs).
Pairs are independent, so N ranks each process their own B pairs with no data-path collective ("scaling": "weak").

  value      W*H*D * B * N * K / t, inputs resident in HBM, device-timed (CUDA events, max over ranks)
  e2e        the same metric through the reference-facing C ABI call `stepth_cuda_depth_dense_batch` with HOST
             (pinned) buffers: host->device copies of every input and device->host copies of every depth map inside
             the timed region
  roofline   the dominant kernel (the ring search): algorithmic bytes per launch / its mean launch duration, measured
             with CUDA events inside the timed region, against MEASURED_PEAKS.json's copy bandwidth
  cpu_baseline  the CPU oracle (a C port of the reference path; the reference is Rust and cannot be built here) timed
             on this box's cores on a bounded sample of the same workload (rank 0, N = 1 only)

`--impl reference` times that CPU port alone (reference-faithful threading: the reference splits the query list in
len/8-sized chunks -> 8-9 rayon tasks, depth_image.rs:111-112) and prints the same JSON line with "impl": "reference".
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

W, H, D = 1920, 1080, 128          # BASELINE.json configs[1] (set per workload in main())
PRECISION = [36, 36, 36]           # u8::MAX / 7 (Readme.md:15)
METRIC = "Mpx*disparities/sec"
FUSED = False                      # invert_depth + select_foreground + apply_mask fused into the step (configs[2..])

# name -> (w, h, max_rings, pairs per step [per GPU | total], fused tail, scaling)
WORKLOADS = {
    "1080p": (1920, 1080, 128, 16, False, "weak"),        # configs[1]: the metric's configuration (default)
    "4k_fused": (3840, 2160, 256, 4, True, "weak"),       # configs[2]
    "batch4k": (3840, 2160, 256, 64, True, "strong"),     # configs[3]: 64 pairs in total, dealt to the ranks
    "giant": (16384, 8192, 256, 1, False, "strong"),      # configs[4]: one pair, row-banded over the ranks
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default="nat", choices=["nat", "geo", "worst"])
    ap.add_argument("--workload", default="1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--pairs", type=int, default=0, help="pairs per step per GPU (default: the workload's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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


def ncu_traffic(workload, kind):
    """dram bytes (read + write) of the search-stage kernels per launch sequence, from the committed ncu capture"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(f"{workload}_{kind}")
    return None


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while the timed region runs"""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
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
            self._stop_evt.wait(0.004)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path, timed on this box's cores (never the product path)
# ---------------------------------------------------------------------------------------------------------------
def cpu_pass(oracle, m, a, threads):
    """one whole reference pass (depth_image.rs:96-136) over one pair on the CPU"""
    return oracle.depth_dense(m, a, PRECISION, D, resample=oracle.RESAMPLE_U8, threads=threads)


def cpu_baseline(kind: str, budget_s: float = 20.0):
    """bounded sample: whole pairs of the bench workload, as many as fit ~budget_s (at least one)"""
    from oracle import oracle
    from stepth_b200 import synth

    cores = os.cpu_count() or 1
    threads = cores                      # all-cores variant (one chunk per hardware thread)
    done, t_total = 0, 0.0
    while done < 4 and t_total < budget_s:
        m, a = synth.make_pair(kind, W, H, seed=done + 1)
        m, a = m.numpy(), a.numpy()
        t0 = time.perf_counter()
        cpu_pass(oracle, m, a, threads)
        t_total += time.perf_counter() - t0
        done += 1
    val = done * W * H * D / t_total / 1e6
    return {"value": round(val, 3), "unit": METRIC, "cores": threads, "kind": "port",
            "sample": f"{done} whole {W}x{H} '{kind}' pair(s), D={D}, oracle depth_dense with {threads} threads, {t_total:.1f} s"}


def algorithmic_probes(kind: str):
    """probes the reference's scan order needs per query (every pixel of a ring counted once), counted
    by the oracle on pair seed 1 of the bench batch (the pairs of a batch are draws of the same distribution)"""
    from oracle import oracle
    from stepth_b200 import synth

    m, a = synth.make_pair(kind, W, H, seed=1)
    oracle.search_dense(m.numpy(), a.numpy(), PRECISION, D, threads=os.cpu_count() or 1, want_probes=True)
    return oracle.last_unique_probes() / float(W * H)


def run_reference(args):
    """--impl reference: the CPU port with the reference's own parallelism (8-9 chunks), each step one pair"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    from stepth_b200 import synth

    cores = os.cpu_count() or 1
    pairs = []
    for i in range(min(args.steps + args.warmup, 4)):
        m, a = synth.make_pair(args.kind, W, H, seed=i + 1)
        pairs.append((m.numpy(), a.numpy()))
    for i in range(args.warmup):
        cpu_pass(oracle, *pairs[i % len(pairs)], 0)
    t0 = time.perf_counter()
    for i in range(args.steps):
        cpu_pass(oracle, *pairs[i % len(pairs)], 0)   # threads = 0 -> len/8-sized chunks = 8-9 threads (reference)
    dt = time.perf_counter() - t0
    val = args.steps * W * H * D / dt / 1e6
    used = min(9, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": METRIC, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"{args.workload}_d{D}_{args.kind}", "w": W, "h": H, "max_rings": D, "precision": PRECISION,
                   "pairs_per_step": 1, "note": "each step = one pair of the GPU arm's batch (bounded sample)"},
        "cpu_baseline": {"value": round(val, 3), "unit": METRIC, "cores": used, "kind": "port",
                         "sample": f"{args.steps} steps x 1 pair, reference-faithful 8-9 chunk threads on {cores} cores"},
        "e2e": {"value": round(val, 3), "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def time_resident(torch, dist, step, steps, warmup, world):
    """K timed steps over HBM-resident inputs; `step(events)` enqueues one step and records (start, end) CUDA events
    around its ring-search stage.  Returns (total ms [max over ranks], search-stage ms summed over the steps)."""
    for _ in range(warmup):
        step(None)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0.record()
    for i in range(steps):
        step(ev[i])
    t1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total = t0.elapsed_time(t1)
    search = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        t = torch.tensor([total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total = float(t.item())
    return total, search


def time_e2e(torch, dist, step, steps, warmup, world):
    """`step()` = the C-ABI host-buffer call: H2D of all inputs + the whole path + D2H of the results (synchronous)"""
    for _ in range(max(1, min(warmup, 2))):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return dt


def run_ours(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- stepth_b200 has no CPU fallback (use --impl reference for the CPU port)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import stepth_b200 as sb
    from stepth_b200 import bands, device, synth

    ctx = sb.Context(local)
    dev = torch.device("cuda", local)
    scaling = WORKLOADS[args.workload][5]
    flags = sb.RESAMPLE_U8 | ((sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY) if FUSED else 0)
    if args.workload == "giant":
        return run_giant(args, torch, dist, ctx, dev, rank, world, flags)

    # pairs this rank processes per step: weak scaling = a fixed batch per GPU; batch4k = 64 pairs dealt to the ranks
    if scaling == "strong":
        my_pairs = bands.shard_pairs(args.pairs, rank, world)
        total_pairs = args.pairs
    else:
        my_pairs = [rank * args.pairs + i for i in range(args.pairs)]
        total_pairs = args.pairs * world
    B = len(my_pairs)
    SUB = min(B, 16)                                   # pairs per launch sequence (bounds the scratch in HBM)

    def make_batch(kind):
        ms, as_ = [], []
        for i in my_pairs:
            m, a = synth.make_pair(kind, W, H, seed=i + 1, device=dev)
            ms.append(m)
            as_.append(a)
        return torch.stack(ms).contiguous(), torch.stack(as_).contiguous()

    mains, adds = make_batch(args.kind)
    pipes = {}

    def pipe_for(n):
        if n not in pipes:
            pipes[n] = device.DensePipeline(ctx, n, W, H, W, H, PRECISION, D, flags)
        return pipes[n]

    def make_step(mains, adds):
        def step(events):
            for j, i0 in enumerate(range(0, B, SUB)):
                n = min(SUB, B - i0)
                # the search-stage events bracket the first launch sequence of the step (all are alike)
                pipe_for(n).run(mains[i0:i0 + n], adds[i0:i0 + n], search_events=events if j == 0 else None)
        return step

    step = make_step(mains, adds)
    sampler = ClockSampler(local)
    l0 = ctx.launch_count()
    step(None)
    torch.cuda.synchronize()
    launches_per_step = ctx.launch_count() - l0
    for p in pipes.values():
        assert int(p.status.abs().sum()) == 0, "a pair of the bench batch failed (max == 0?)"

    sampler.start()   # samples through the resident AND the end-to-end timed regions (stopped below)
    total_ms, search_ms = time_resident(torch, dist, step, args.steps, args.warmup, world)
    px_disp = W * H * D
    value = total_pairs * args.steps * px_disp / (total_ms * 1e-3) / 1e6

    hbm, which = peaks()
    search_launch_ms = search_ms / args.steps
    alg = search_alg_bytes(W, H, D, SUB)
    achieved = alg / (search_launch_ms * 1e-3) / 1e9
    n_seq = (B + SUB - 1) // SUB
    roofline = {"bound": "hbm", "achieved": round(achieved, 2), "peak": hbm, "unit": "GB/s",
                "frac": round(achieved / hbm, 5), "traffic": ncu_traffic(args.workload, args.kind), "peak_source": which,
                "kernel": "ring search stage (stepth_cuda_dev_search_dense: tile_planes + coarse_transpose + search_near + search_far)",
                "alg_bytes_per_launch": alg, "pairs_per_launch": SUB, "launch_ms": round(search_launch_ms, 4),
                "share_of_step": round(search_ms * n_seq / total_ms, 4),
                "note": "the stage is issue/latency bound, not HBM bound (DESIGN.md section 5); frac is reported as the contract asks"}

    e2e = None
    if not args.no_e2e:
        host_m = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(B)]
        host_a = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(B)]
        out = {"depth": [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)]}
        if FUSED:
            out["mask"] = [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)]
            out["rgba"] = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(B)]
        for i in range(B):
            host_m[i].copy_(mains[i])
            host_a[i].copy_(adds[i])

        def e2e_step():
            for i0 in range(0, B, SUB):
                sl = slice(i0, min(B, i0 + SUB))
                ctx.depth_dense_batch(host_m[sl], host_a[sl], PRECISION, D, flags, out={k: v[sl] for k, v in out.items()})

        e2e_steps = max(2, min(args.steps, 5))
        e2e_ms = time_e2e(torch, dist, e2e_step, e2e_steps, args.warmup, world)
        rw, rh = reach(W, H, D)
        e2e = {"value": round(total_pairs * e2e_steps * px_disp / (e2e_ms * 1e-3) / 1e6, 3), "unit": METRIC,
               "h2d_bytes_per_step": B * (4 * W * H + 3 * rw * rh), "d2h_bytes_per_step": B * W * H * (6 if FUSED else 1),
               "ms_per_step": round(e2e_ms / e2e_steps, 3), "steps": e2e_steps,
               "api": "stepth_cuda_depth_dense_batch (host buffers, pinned; chunks pipelined over copy/compute/copy streams)"}
        # the resident pipeline and the host-buffer call must agree bit for bit
        assert torch.equal(out["depth"][0], pipe_for(min(SUB, B)).depth[0].cpu()) or B > SUB, "host-buffer path and device-resident path disagree"

    clocks = sampler.stop()
    extra = {}
    if args.extra_kinds and world == 1:
        for kind in ("geo", "worst"):
            if kind == args.kind:
                continue
            m2, a2 = make_batch(kind)
            steps2 = max(1, min(args.steps, 3))
            t_ms, s_ms = time_resident(torch, dist, make_step(m2, a2), steps2, 1, world)
            extra[kind] = {"value": round(total_pairs * steps2 * px_disp / (t_ms * 1e-3) / 1e6, 3), "ms_per_step": round(t_ms / steps2, 3),
                           "search_ms_per_launch": round(s_ms / steps2, 3)}
            del m2, a2

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args.kind)
        # the other half of SURVEY 8(d)'s roofline, t_roof = max(B_alg / BW_HBM, O_alg / P_INT): O_alg = 2 integer ops
        # per probe of the reference order (oracle-counted), P_INT measured on this GPU by a register-only kernel
        probes_q = algorithmic_probes(args.kind)
        p_mixed, p_logic = ctx.int_peak(False), ctx.int_peak(True)
        ops = 2.0 * probes_q * W * H * SUB
        t_int_ms, t_hbm_ms = ops / p_mixed * 1e3, alg / (hbm * 1e9) * 1e3
        roofline["int"] = {
            "bound": "int-alu", "unit": "Tiop/s", "ops_per_launch": round(ops), "probes_per_query": round(probes_q, 2),
            "achieved": round(ops / (search_launch_ms * 1e-3) / 1e12, 3), "peak": round(p_mixed / 1e12, 3),
            "frac": round(ops / (search_launch_ms * 1e-3) / p_mixed, 4), "peak_logic_only": round(p_logic / 1e12, 3),
            "peak_source": "measured (stepth_cuda_int_peak: LOP3+IMAD.IADD on both integer pipes; logic_only = LOP3+SHF, alu pipe)",
            "t_roof_ms": round(max(t_int_ms, t_hbm_ms), 4), "t_hbm_ms": round(t_hbm_ms, 4), "t_int_ms": round(t_int_ms, 4),
            "frac_of_t_roof": round(max(t_int_ms, t_hbm_ms) / search_launch_ms, 4),
            "note": "O_alg = 2 ops x probes the reference order needs (oracle count, seed-1 pair x pairs per launch); "
                    "the tile search skips probes, so frac may exceed 1 (SURVEY.md 8(d))"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"{args.workload}_d{D}_{args.kind}", "w": W, "h": H, "max_rings": D, "precision": PRECISION,
                       "pairs_per_step_total": total_pairs, "pairs_per_step_this_gpu": B,
                       "flags": "RESAMPLE_U8" + ("|INVERT|FOREGROUND|APPLY" if FUSED else ""),
                       "l2": f"inputs {B * (4 + 3) * W * H / 1e6:.0f} MB per step per GPU > 126 MB L2 (no flush needed)",
                       "parallelism": f"pair-batch x{world} (no collective)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
        }
        if extra:
            line["other_distributions"] = extra
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_giant(args, torch, dist, ctx, dev, rank, world, flags):
    """configs[4]: ONE pair, main rows banded over the ranks (bands.BandRunner): MAX all-reduce of the normaliser, 3-row
    halo exchange, and an NCCL all-gather of the depth bands so that every GPU ends with the assembled map."""
    import stepth_b200 as sb
    from stepth_b200 import bands, synth

    runner = bands.BandRunner(bands.DeviceStages(ctx), W, H, H, PRECISION, D, flags)
    b = runner.band
    # every rank generates the same pair from the same seed and keeps only its rows (synthetic data; no scatter timed)
    m, a = synth.make_pair(args.kind, W, H, seed=1, device=dev)
    mb, ab = m[None, b.y0:b.y1].contiguous(), a[None, b.ay0:b.ay1].contiguous()
    del m, a
    torch.cuda.empty_cache()
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    l0 = ctx.launch_count()
    out = runner.run(mb, ab, gather_depth=world > 1)
    torch.cuda.synchronize()
    launches = ctx.launch_count() - l0
    assert out["status"] == 0

    def step(events):
        if events is not None:
            events[0].record()
        runner.run(mb, ab, gather_depth=world > 1)
        if events is not None:
            events[1].record()

    sampler.start()
    total_ms, _ = time_resident(torch, dist, step, args.steps, args.warmup, world)
    clocks = sampler.stop()
    value = args.steps * W * H * D / (total_ms * 1e-3) / 1e6
    # one more (untimed) run with CUDA events between the stages: where each rank's time goes
    runner.profile = True
    runner.run(mb, ab, gather_depth=world > 1)
    runner.profile = False
    stage_ms = [runner.stage_ms]
    if world > 1:
        stage_ms = [None] * world
        dist.all_gather_object(stage_ms, runner.stage_ms)
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"giant_d{D}_{args.kind}", "w": W, "h": H, "max_rings": D, "precision": PRECISION,
                       "bands": [[p.y0, p.y1, p.ay0, p.ay1] for p in runner.plan],
                       "l2": "inputs larger than L2",
                       "parallelism": f"row bands x{world}: allreduce(MAX) + 3-row halo all-gather + NCCL all-gather of depth bands"},
            "roofline": None, "cpu_baseline": None, "e2e": None, "gpu_launches": launches * args.steps, "clocks": clocks,
            "stage_ms_per_rank": stage_ms,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    global W, H, D, FUSED
    args = parse_args()
    w, h, d, pairs, fused, _ = WORKLOADS[args.workload]
    W, H, D, FUSED = w, h, d, fused
    if not args.pairs:
        args.pairs = pairs
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
