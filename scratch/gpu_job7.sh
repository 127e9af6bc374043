#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --clock-control none --import-source on -k regex:"search_" -c 3 -f -o gpurun_out/src_prof python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/src_prof.log 2>&1; echo "rc=$?"
