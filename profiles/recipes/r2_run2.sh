#!/bin/bash
# round 2, second GPU call: staging / RGB tests, default bench with the pinned ring, link-ceiling probe, ncu captures of HEAD
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "pageable or rgb_main or apply_needs or without_reachable or probe_counter" > gpurun_out/r2_pytest_new.log 2>&1; echo "pytest(new) rc=$?"; tail -5 gpurun_out/r2_pytest_new.log
timeout 300 python profiles/recipes/h2d_ceiling.py > gpurun_out/r2_h2d_ceiling_n1.json 2>gpurun_out/r2_h2d.err; echo "h2d rc=$?"; cat gpurun_out/r2_h2d_ceiling_n1.json
for t in 2 4 8; do STEPTH_STAGE_THREADS=$t timeout 300 python bench.py --steps 4 --warmup 3 --no-others --no-cpu-baseline --no-parity > gpurun_out/r2_bench_stage_t$t.json 2>gpurun_out/r2_bench_stage_t$t.err; echo "stage threads $t rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_stage_t$t.json')); print(l['e2e']['ms_per_step'], l['e2e']['pageable'])"; done
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_default.err
# ncu: launch list of two steps of the 1080p workload, then full captures of one launch sequence (1080p, then 4K)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_|mask_apply|depth_split" -c 60 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 6 -f -o gpurun_out/r2_1080p_full python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full 1080p rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"search_|tile_planes|coarse_transpose" -c 4 -f -o gpurun_out/r2_batch4k_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full4k.log 2>&1; echo "ncu full 4k rc=$?"
ls -la gpurun_out/*.ncu-rep
