#!/usr/bin/env python
"""The end-to-end roofline of the batch path: what the host<->device links of this box carry with NO kernels at all.

    python profiles/recipes/h2d_ceiling.py                                  (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        profiles/recipes/h2d_ceiling.py                                     (N processes, one per GPU, all at once)

Every rank moves, per iteration, what one rank's share of a batch4k step moves (64 / N pairs of 3840x2160: 41.9 MB up and
49.8 MB down per pair) between page-locked host memory and its GPU, uploads and downloads on two streams at the same time,
in 32 MiB copies.  Prints one JSON line: aggregate and per-GPU GB/s for up, down and both together."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    bound = None
    try:
        import pynvml

        pynvml.nvmlInit()
        if os.environ.get("NO_BIND") != "1":
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        bound = len(os.sched_getaffinity(0))
    except Exception:
        pass
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pairs = max(1, 64 // world)
    up, down = pairs * 41_899_008, pairs * 49_766_400
    piece = 32 << 20
    h_in = torch.empty(min(up, 1 << 30), dtype=torch.uint8).pin_memory()
    h_out = torch.empty(min(down, 1 << 30), dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty_like(h_in, device="cuda"), torch.empty_like(h_out, device="cuda")
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

    def run(do_up, do_down):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if do_up:
            with torch.cuda.stream(s_up):
                for o in range(0, up, piece):
                    a = o % h_in.numel()
                    n = min(piece, h_in.numel() - a, up - o)
                    d_in[a:a + n].copy_(h_in[a:a + n], non_blocking=True)
        if do_down:
            with torch.cuda.stream(s_down):
                for o in range(0, down, piece):
                    a = o % h_out.numel()
                    n = min(piece, h_out.numel() - a, down - o)
                    h_out[a:a + n].copy_(d_out[a:a + n], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt

    res = {}
    for name, (u, d) in {"up": (True, False), "down": (False, True), "both": (True, True)}.items():
        run(u, d)
        best = min(run(u, d) for _ in range(4))
        moved = (up if u else 0) + (down if d else 0)
        res[name] = {"ms": round(best * 1e3, 3), "gbs_per_gpu": round(moved / best / 1e9, 2), "gbs_aggregate": round(moved * world / best / 1e9, 2)}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "pairs_per_gpu": pairs, "up_bytes_per_gpu": up, "down_bytes_per_gpu": down,
                          "cpus": os.cpu_count(), "cpus_bound": bound, **res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
