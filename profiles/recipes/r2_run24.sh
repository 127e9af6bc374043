#!/bin/bash
# round 2, twenty-fourth GPU call: task slots per warp of the origin-tile stage (shared memory per CTA -> L1 carve-out)
mkdir -p gpurun_out
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for wl in 1080p 4k_fused; do
run t256 $wl X=1
run t384 $wl STEPTH_LIB=stepth_b200/libstepth_cuda_t384.so
run t512 $wl STEPTH_LIB=stepth_b200/libstepth_cuda_t512.so
done
