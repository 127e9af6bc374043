#!/bin/bash
# round 2, tenth GPU call: origin-tile near stage -- CTA size x register cap x rings walked
mkdir -p gpurun_out
B="python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for v in w4mb10 w2mb16 w2mb20 w2mb21 w1mb32 w8mb5; do
  for r in 6 12; do run ${v}_r$r STEPTH_LIB=stepth_b200/libstepth_cuda_$v.so STEPTH_OT_RINGS=$r; done
done
run w2mb20_r16 STEPTH_LIB=stepth_b200/libstepth_cuda_w2mb20.so STEPTH_OT_RINGS=16
run w2mb20_r12_far8 STEPTH_LIB=stepth_b200/libstepth_cuda_w2mb20.so STEPTH_OT_RINGS=12
