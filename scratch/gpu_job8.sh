#!/bin/bash
# 8-GPU box: scaling of the three multi-GPU workloads
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_1080p_n8.json 2> gpurun_out/b8a.err; echo "1080p n8 rc=$?"; tail -2 gpurun_out/b8a.err
grep '^{' gpurun_out/bench_1080p_n8.json | cut -c1-330
timeout 600 $TR bench.py --gpus 8 --workload batch4k --steps 3 --warmup 3 > gpurun_out/bench_batch4k_n8.json 2> gpurun_out/b8b.err; echo "batch4k n8 rc=$?"; tail -2 gpurun_out/b8b.err
grep '^{' gpurun_out/bench_batch4k_n8.json | cut -c1-330
timeout 600 $TR bench.py --gpus 8 --workload giant --steps 3 --warmup 3 > gpurun_out/bench_giant_n8.json 2> gpurun_out/b8c.err; echo "giant n8 rc=$?"; tail -2 gpurun_out/b8c.err
grep '^{' gpurun_out/bench_giant_n8.json | cut -c1-330
timeout 600 python bench.py --workload batch4k --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_batch4k_n1.json 2> gpurun_out/b1b.err; echo "batch4k n1 rc=$?"
grep '^{' gpurun_out/bench_batch4k_n1.json | cut -c1-330
