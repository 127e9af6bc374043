#!/bin/bash
# round 2, seventh GPU call (2 GPUs): the multi-GPU entry points below the boundary, bench at N=2, link ceiling at N=2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_n2.txt 2>&1; nproc >> gpurun_out/r2_topo_n2.txt
timeout 900 python -m pytest tests -m gpu -x -q -k "multi or nccl or comm_band or band" > gpurun_out/r2_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest_n2.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $T profiles/recipes/h2d_ceiling.py > gpurun_out/r2_h2d_ceiling_n2.json 2>gpurun_out/r2_h2d_n2.err; echo "h2d rc=$?"; cat gpurun_out/r2_h2d_ceiling_n2.json
timeout 600 $T bench.py --gpus 2 --workload giant --steps 10 --warmup 3 --no-others > gpurun_out/r2_bench_giant_n2.json 2>gpurun_out/r2_bench_giant_n2.err; echo "giant n2 rc=$?"; tail -3 gpurun_out/r2_bench_giant_n2.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_giant_n2.json')); print(l['ms_per_step'], l['e2e'], l['parity'], l['stage_ms_per_rank'])"
timeout 900 $T bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_default_n2.json 2>gpurun_out/r2_bench_default_n2.err; echo "default n2 rc=$?"; tail -3 gpurun_out/r2_bench_default_n2.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_default_n2.json')); print(l['ms_per_step'], l['e2e'], l['parity']); print({k:(v['ms_per_step'], v['e2e']['ms_per_step'] if v['e2e'] else None, v['parity']['bit_exact']) for k,v in l['other_workloads'].items()})"
