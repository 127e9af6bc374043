#!/bin/bash
# round 2, ninth GPU call: the origin-tile near stage (search_otile.cuh) -- whole GPU test-suite, then 1080p lines for the knobs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest_gpu.log
B="python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
run walk STEPTH_NEAR_TILE=0
run otile STEPTH_NEAR_TILE=2
for r in 3 4 8 10 12 16; do run otile_r$r STEPTH_OT_RINGS=$r; done
run otile_half STEPTH_OT_HALVES=1
for mb in 10 12; do run otile_mb$mb STEPTH_LIB=stepth_b200/libstepth_cuda_otmb$mb.so; done
run otile_geo STEPTH_NEAR_TILE=2 BENCH_KIND=geo
timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e --extra-kinds > gpurun_out/r2_v_kinds.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/r2_v_kinds.json')); print(l['ms_per_step'], l.get('other_distributions'))"
timeout 300 python bench.py --workload 1080p --pairs 1 --steps 30 --warmup 5 --no-others --no-cpu-baseline > gpurun_out/r2_v_onepair.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/r2_v_onepair.json')); print('one pair', l['ms_per_step'], l['e2e']['ms_per_step'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 12 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"; grep -v "^==" gpurun_out/r2_launches_1080p.csv | cut -d, -f5,15 | head -8
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_" -c 2 -f -o gpurun_out/r2_otile_full python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
