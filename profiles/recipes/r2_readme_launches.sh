#!/bin/bash
# configs[0] under ncu: the launch list of ONE README pair in block mode (device discretiser D0, K0, K1a' for block lists, K1b / K1c,
# paint, K2) and in dense mode (eighth-tile origin stage, K1c) -- per-launch durations, cold-cache and serialised
mkdir -p gpurun_out
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_readme_blocks.csv python bench.py --workload readme_blocks --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_readme_blocks.log 2>&1; echo "ncu blocks rc=$?"
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2_launches_readme_dense.csv python bench.py --workload readme_dense --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_readme_dense.log 2>&1; echo "ncu dense rc=$?"
wc -l gpurun_out/r2_launches_readme_*.csv
