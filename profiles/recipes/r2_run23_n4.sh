#!/bin/bash
# round 2, twenty-third GPU call (4 GPUs): the default bench line at N=4 and N=2 (scaling table from HEAD)
mkdir -p gpurun_out
for n in 4 2; do
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533"
timeout 1200 $T bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r2_bench_default_n$n.json 2>gpurun_out/r2_bench_default_n$n.err; echo "default n$n rc=$?"; tail -2 gpurun_out/r2_bench_default_n$n.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_default_n$n.json')); print(l['ms_per_step'], l['value'], l['e2e']['ms_per_step'], l['e2e'].get('frac_of_link_ceiling'), l['parity']['bit_exact']); print({k:(v['ms_per_step'], v['e2e']['ms_per_step'] if v['e2e'] else None, v['parity']['bit_exact']) for k,v in l['other_workloads'].items()}); print(l['other_workloads']['giant']['stage_ms_per_rank'])"
done
