#!/bin/bash
# round 2, fifteenth GPU call (8 GPUs): the multi-GPU tests on all 8 GPUs, the link ceiling, giant over peer windows and over NCCL,
# the default bench line
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_n8.txt 2>&1; nproc >> gpurun_out/r2_topo_n8.txt
timeout 1200 python -m pytest tests -m gpu -x -q -k "multi or nccl or comm_band or band" > gpurun_out/r2_pytest_n8.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest_n8.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $T profiles/recipes/h2d_ceiling.py > gpurun_out/r2_h2d_ceiling_n8.json 2>gpurun_out/r2_h2d_n8.err; echo "h2d rc=$?"; tail -2 gpurun_out/r2_h2d_ceiling_n8.json
for peer in 1 0; do
STEPTH_PEER=$peer timeout 600 $T bench.py --gpus 8 --workload giant --steps 20 --warmup 5 --no-others > gpurun_out/r2_bench_giant_n8_peer$peer.json 2>gpurun_out/r2_bench_giant_n8.err; echo "giant n8 peer=$peer rc=$?"; tail -3 gpurun_out/r2_bench_giant_n8.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_giant_n8_peer$peer.json')); print(l['ms_per_step'], l['e2e']['ms_per_step'], l['parity']['bit_exact'], l['config']['exchange'][:12]); [print('  ', s) for s in l['stage_ms_per_rank']]"
done
timeout 1200 $T bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2_bench_default_n8.json 2>gpurun_out/r2_bench_default_n8.err; echo "default n8 rc=$?"; tail -3 gpurun_out/r2_bench_default_n8.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_default_n8.json')); print(l['ms_per_step'], l['value'], l['e2e'], l['parity']['bit_exact']); print({k:(v['ms_per_step'], v['e2e']['ms_per_step'] if v['e2e'] else None, v['parity']['bit_exact']) for k,v in l['other_workloads'].items()})"
