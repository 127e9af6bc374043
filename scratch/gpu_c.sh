#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
for nr in 6 5 8; do
  STEPTH_NEAR_RINGS=$nr timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --extra-kinds > gpurun_out/b.json 2> gpurun_out/b.err
  python -c "
import json; d=json.load(open('gpurun_out/b.json')); o=d['other_distributions']; print('near=$nr step', d['ms_per_step'], 'search', d['roofline']['launch_ms'], 'geo', o['geo']['search_ms_per_launch'], 'worst', o['worst']['search_ms_per_launch'])"
done
for m in 0 10; do
  STEPTH_MID_RINGS=$m timeout 300 python bench.py --workload 4k_fused --steps 8 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err
  python -c "
import json; d=json.load(open('gpurun_out/b.json')); print('4k mid=$m step', d['ms_per_step'], 'search', d['roofline']['launch_ms'])"
done
