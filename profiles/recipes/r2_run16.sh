#!/bin/bash
# round 2, sixteenth GPU call: rings per pass of the origin-tile stage once a single batch is left
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for wl in 1080p 4k_fused; do
run rpp2 $wl X=1
for r in 1 3 4; do run rpp$r $wl STEPTH_LIB=stepth_b200/libstepth_cuda_rpp$r.so; done
run rpp2_r16 $wl STEPTH_OT_RINGS=16
run rpp4_r16 $wl STEPTH_LIB=stepth_b200/libstepth_cuda_rpp4.so STEPTH_OT_RINGS=16
done
run rpp2_geo 1080p X=1 
