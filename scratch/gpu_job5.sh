#!/bin/bash
# 2-GPU box: all GPU tests (incl. NCCL bands + C++ mirror), giant at N=1/2, default bench at N=2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_L.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -12 gpurun_out/pytest_gpu2.log
timeout 600 python bench.py --workload giant --steps 2 --warmup 3 > gpurun_out/bench_giant_n1.json 2> gpurun_out/benchg.err; echo "giant n1 rc=$?"; tail -3 gpurun_out/benchg.err
cat gpurun_out/bench_giant_n1.json | cut -c1-600
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --workload giant --steps 2 --warmup 3 > gpurun_out/bench_giant_n2.json 2> gpurun_out/benchg2.err; echo "giant n2 rc=$?"; tail -3 gpurun_out/benchg2.err
cat gpurun_out/bench_giant_n2.json | cut -c1-600
timeout 600 $TR bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_1080p_n2.json 2> gpurun_out/bench2.err; echo "1080p n2 rc=$?"; tail -3 gpurun_out/bench2.err
cat gpurun_out/bench_1080p_n2.json | cut -c1-400
timeout 600 $TR bench.py --gpus 2 --workload batch4k --steps 2 --warmup 3 --no-e2e > gpurun_out/bench_batch4k_n2.json 2> gpurun_out/bench3.err; echo "batch4k n2 rc=$?"; tail -3 gpurun_out/bench3.err
cat gpurun_out/bench_batch4k_n2.json | cut -c1-400
