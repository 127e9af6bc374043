#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"finish_" -c 2 -f -o gpurun_out/fin2_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_fin.log 2>&1; echo "ncu fin rc=$?"
