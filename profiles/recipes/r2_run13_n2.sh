#!/bin/bash
# round 2, thirteenth GPU call (2 GPUs): the band exchanges over peer windows (csrc/peer.cuh) against NCCL: tests, giant bench lines
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_n2.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q -k "multi or nccl or comm_band or band" > gpurun_out/r2_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2_pytest_n2.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
for peer in 1 0; do
STEPTH_PEER=$peer timeout 600 $T bench.py --gpus 2 --workload giant --steps 10 --warmup 3 --no-others > gpurun_out/r2_bench_giant_n2_peer$peer.json 2>gpurun_out/r2_bench_giant_n2.err; echo "giant n2 peer=$peer rc=$?"; tail -3 gpurun_out/r2_bench_giant_n2.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_giant_n2_peer$peer.json')); print(l['ms_per_step'], l['e2e']['ms_per_step'], l['parity']['bit_exact'], l['config']['exchange'][:12]); [print('  ', s) for s in l['stage_ms_per_rank']]"
done
timeout 600 python bench.py --workload giant --steps 10 --warmup 3 --no-others --no-cpu-baseline > gpurun_out/r2_bench_giant_n1.json 2>gpurun_out/r2_bench_giant_n1.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_giant_n1.json')); print('n1', l['ms_per_step'], l['e2e']['ms_per_step'], l['parity']['bit_exact']); [print('  ', s) for s in l['stage_ms_per_rank']]"
