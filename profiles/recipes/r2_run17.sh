#!/bin/bash
# round 2, seventeenth GPU call: A/B of the near stage before / after the rings-per-pass restructuring on ONE box (boxes differ by a few percent)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.mem,temperature.gpu,power.draw --format=csv > gpurun_out/r2_box.txt
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'])" || tail -3 gpurun_out/r2_v_$name.err; }
for i in 1 2; do
run cur 1080p X=1
run prev 1080p STEPTH_LIB=stepth_b200/libstepth_cuda_prev.so
run walk 1080p STEPTH_NEAR_TILE=0
done
run cur 4k_fused X=1
run prev 4k_fused STEPTH_LIB=stepth_b200/libstepth_cuda_prev.so
