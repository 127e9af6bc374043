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

W, H, D = 1920, 1080, 128          # BASELINE.json configs[1]
PRECISION = [36, 36, 36]           # u8::MAX / 7 (Readme.md:15)
METRIC = "Mpx*disparities/sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default="nat", choices=["nat", "geo", "worst"])
    ap.add_argument("--pairs", type=int, default=16, help="pairs per step per GPU")
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
            self._stop_evt.wait(0.05)

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
        "config": {"workload": f"1080p_d{D}_{args.kind}", "w": W, "h": H, "max_rings": D, "precision": PRECISION,
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
def time_resident(torch, dist, pipe, mains, adds, steps, warmup, world):
    """K timed steps over HBM-resident inputs; returns (total ms [max over ranks], search-stage ms summed)"""
    for _ in range(warmup):
        pipe.run(mains, adds)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0.record()
    for i in range(steps):
        pipe.run(mains, adds, search_events=ev[i])
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


def time_e2e(torch, dist, ctx, sb, host_m, host_a, host_d, steps, warmup, world):
    """the C-ABI host-buffer call: H2D of all inputs + the whole path + D2H of the depth maps, per step"""
    def step():
        ctx.depth_dense_batch(host_m, host_a, PRECISION, D, sb.RESAMPLE_U8, out={"depth": host_d})

    for _ in range(max(1, min(warmup, 2))):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()                       # synchronous: returns after the device->host copies finished
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
    from stepth_b200 import device, synth

    ctx = sb.Context(local)
    B = args.pairs
    dev = torch.device("cuda", local)

    def make_batch(kind):
        ms, as_ = [], []
        for i in range(B):
            m, a = synth.make_pair(kind, W, H, seed=rank * B + i + 1, device=dev)
            ms.append(m)
            as_.append(a)
        return torch.stack(ms).contiguous(), torch.stack(as_).contiguous()

    mains, adds = make_batch(args.kind)
    flags = sb.RESAMPLE_U8
    pipe = device.DensePipeline(ctx, B, W, H, W, H, PRECISION, D, flags)
    sampler = ClockSampler(local)
    l0 = ctx.launch_count()
    pipe.run(mains, adds)
    torch.cuda.synchronize()
    launches_per_step = ctx.launch_count() - l0
    assert int(pipe.status.abs().sum()) == 0, "a pair of the bench batch failed (max == 0?)"

    sampler.start()
    total_ms, search_ms = time_resident(torch, dist, pipe, mains, adds, args.steps, args.warmup, world)
    clocks = sampler.stop()
    px_disp = W * H * D
    value = world * B * args.steps * px_disp / (total_ms * 1e-3) / 1e6

    hbm, which = peaks()
    search_launch_ms = search_ms / args.steps
    alg = search_alg_bytes(W, H, D, B)
    achieved = alg / (search_launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": round(achieved, 2), "peak": hbm, "unit": "GB/s",
                "frac": round(achieved / hbm, 5), "traffic": None, "peak_source": which,
                "kernel": "ring search stage (stepth_cuda_dev_search_dense)", "alg_bytes_per_launch": alg,
                "launch_ms": round(search_launch_ms, 4), "share_of_step": round(search_ms / total_ms, 4)}

    e2e = None
    if not args.no_e2e:
        host_m = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(B)]
        host_a = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(B)]
        host_d = [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)]
        for i in range(B):
            host_m[i].copy_(mains[i])
            host_a[i].copy_(adds[i])
        e2e_steps = max(2, min(args.steps, 5))
        e2e_ms = time_e2e(torch, dist, ctx, sb, host_m, host_a, host_d, e2e_steps, args.warmup, world)
        rw, rh = reach(W, H, D)
        e2e = {"value": round(world * B * e2e_steps * px_disp / (e2e_ms * 1e-3) / 1e6, 3), "unit": METRIC,
               "h2d_bytes_per_step": B * (4 * W * H + 3 * rw * rh), "d2h_bytes_per_step": B * W * H,
               "ms_per_step": round(e2e_ms / e2e_steps, 3), "steps": e2e_steps,
               "api": "stepth_cuda_depth_dense_batch (host buffers, pinned)"}
        # the resident pipeline and the host-buffer call must agree bit for bit
        assert torch.equal(host_d[0], pipe.depth[0].cpu()), "host-buffer path and device-resident path disagree"

    extra = {}
    if args.extra_kinds and world == 1:
        for kind in ("geo", "worst"):
            if kind == args.kind:
                continue
            m2, a2 = make_batch(kind)
            steps2 = max(1, min(args.steps, 3))
            t_ms, s_ms = time_resident(torch, dist, pipe, m2, a2, steps2, 1, world)
            extra[kind] = {"value": round(B * steps2 * px_disp / (t_ms * 1e-3) / 1e6, 3), "ms_per_step": round(t_ms / steps2, 3),
                           "search_ms_per_step": round(s_ms / steps2, 3)}
            del m2, a2

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args.kind)

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"1080p_d{D}_{args.kind}", "w": W, "h": H, "max_rings": D, "precision": PRECISION,
                       "pairs_per_step_per_gpu": B, "flags": "RESAMPLE_U8",
                       "l2": f"inputs {B * (4 + 3) * W * H / 1e6:.0f} MB per step > 126 MB L2 (no flush needed)",
                       "parallelism": f"pair-batch x{world} (no collective)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
        }
        if extra:
            line["other_distributions"] = extra
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
