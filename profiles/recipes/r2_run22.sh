#!/bin/bash
# round 2, twenty-second GPU call: origin-tile stage with fan-out (idle lanes of a short list walk the following rings): parity, A/B
mkdir -p gpurun_out
STEPTH_LIB=stepth_b200/libstepth_cuda_fan.so timeout 1500 python -m pytest tests -m gpu -x -q -k "not cpp_host" > gpurun_out/r2_pytest_fan.log 2>&1; echo "pytest(fan) rc=$?"; tail -3 gpurun_out/r2_pytest_fan.log
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
for i in 1 2; do
run cur 1080p X=1
run fan 1080p STEPTH_LIB=stepth_b200/libstepth_cuda_fan.so
done
run cur 4k_fused X=1
run fan 4k_fused STEPTH_LIB=stepth_b200/libstepth_cuda_fan.so
run fan_r16 4k_fused STEPTH_LIB=stepth_b200/libstepth_cuda_fan.so STEPTH_OT_RINGS=16
run fan_r16 1080p STEPTH_LIB=stepth_b200/libstepth_cuda_fan.so STEPTH_OT_RINGS=16
