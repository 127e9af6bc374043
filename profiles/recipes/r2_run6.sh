#!/bin/bash
# round 2, sixth GPU call: the TMA-staged tile stage (search_tile.cuh): the whole GPU test-suite, then stage boundaries / occupancy
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest_gpu.log
B="python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])"; }
run walk STEPTH_NEAR_TILE=0
run tile_nr6 STEPTH_NEAR_TILE=1
run tile_nr2 STEPTH_NEAR_RINGS=2
run tile_nr3 STEPTH_NEAR_RINGS=3
run tile_nr4 STEPTH_NEAR_RINGS=4
run tile_nr4_m8 STEPTH_NEAR_RINGS=4 STEPTH_MID_RINGS=8
run tile_nr8 STEPTH_NEAR_RINGS=8
for mb in 3 5 6; do run tile_mb$mb STEPTH_LIB=stepth_b200/libstepth_cuda_tmb$mb.so; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 14 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
