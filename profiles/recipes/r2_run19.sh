#!/bin/bash
# round 2, nineteenth GPU call: origin rows per warp (whole / half / quarter tiles) against the size of the launch; 1- and 3-warp CTAs
mkdir -p gpurun_out
run() { name=$1; wl=$2; pairs=$3; shift; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --pairs $pairs --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', $pairs, l['ms_per_step'], l['roofline']['launch_ms'])" || tail -3 gpurun_out/r2_v_$name.err; }
for p in 2 4 8 16; do for nit in 8 4 2; do run nit${nit} 1080p $p STEPTH_OT_NIT=$nit; done; done
for nit in 8 4; do run nit${nit} 4k_fused 4 STEPTH_OT_NIT=$nit; done
for v in w1 w3; do run $v 1080p 16 STEPTH_LIB=stepth_b200/libstepth_cuda_$v.so; done
run cur 1080p 16 X=1
