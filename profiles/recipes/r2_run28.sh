#!/bin/bash
# round 2, twenty-eighth GPU call: the tail kernel (one warp per query for a short work list): the whole GPU suite, small launches with / without
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest_gpu.log
run() { name=$1; wl=$2; pairs=$3; shift; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --pairs $pairs --steps 30 --warmup 5 --no-others --no-cpu-baseline > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', $pairs, l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for t in 1 0; do
run tail$t readme_dense 1 STEPTH_TAIL=$t
run tail$t readme_blocks 1 STEPTH_TAIL=$t
run tail$t 1080p 1 STEPTH_TAIL=$t
run tail$t 4k_fused 1 STEPTH_TAIL=$t
run tail$t 1080p 16 STEPTH_TAIL=$t
done
