#!/bin/bash
# round 2, fifth GPU call: launch-sequence size (planes resident in L2?), per-line profile of the near kernel at HEAD
mkdir -p gpurun_out
for sub in 2 4 8 16; do STEPTH_BENCH_SUB=$sub timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_sub$sub.json 2>gpurun_out/r2_sub.err; python -c "
import json; l=json.load(open('gpurun_out/r2_sub$sub.json')); print('SUB $sub', l['ms_per_step'], l['roofline']['launch_ms'])"; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_near|search_far" -c 2 -f -o gpurun_out/r2_k1_full python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
