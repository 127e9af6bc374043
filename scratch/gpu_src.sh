#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_near" -c 1 -f -o gpurun_out/near_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_near.log 2>&1; echo "ncu near rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_far" -c 2 -f -o gpurun_out/far_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_far.log 2>&1; echo "ncu far rc=$?"
ls -la gpurun_out
