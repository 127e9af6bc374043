#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tile_planes" -c 1 -f -o gpurun_out/k0_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_k0.log 2>&1; echo "ncu k0 rc=$?"
