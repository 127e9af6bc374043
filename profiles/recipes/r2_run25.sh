#!/bin/bash
# round 2, twenty-fifth GPU call: fine_test with the candidate word first (3 loads, then 12) against all 15 at once
mkdir -p gpurun_out
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for wl in 1080p 4k_fused; do
run all15 $wl X=1
run cand $wl STEPTH_LIB=stepth_b200/libstepth_cuda_cand.so
done
run cand_geo 1080p STEPTH_LIB=stepth_b200/libstepth_cuda_cand.so
timeout 300 python bench.py --workload 1080p --kind geo --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_g0.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/r2_v_g0.json')); print('geo all15', l['ms_per_step'], l['parity']['bit_exact'])"
STEPTH_LIB=stepth_b200/libstepth_cuda_cand.so timeout 300 python bench.py --workload 1080p --kind geo --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_g1.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/r2_v_g1.json')); print('geo cand', l['ms_per_step'], l['parity']['bit_exact'])"
