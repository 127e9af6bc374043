#!/bin/bash
# round 2, eighteenth GPU call: the origin-tile stage with CTAs that own a span of consecutive tiles (warps pull tiles one by one)
mkdir -p gpurun_out
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
run cur 1080p X=1
for v in w4s16 w4s32 w8s32 w2s8 w2s16; do run $v 1080p STEPTH_LIB=stepth_b200/libstepth_cuda_$v.so; done
run cur 1080p X=1
run cur 4k_fused X=1
for v in w4s16 w4s32 w8s32 w2s16; do run $v 4k_fused STEPTH_LIB=stepth_b200/libstepth_cuda_$v.so; done
