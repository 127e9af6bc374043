#!/bin/bash
# round 2, fourth GPU call: rings 0+1 from three row windows + the exact stopping rule (K1), non-temporal staging copies
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest_gpu.log
for nt in 1 0; do for t in 4 8 12; do STEPTH_STAGE_NT=$nt STEPTH_STAGE_THREADS=$t timeout 300 python bench.py --steps 4 --warmup 3 --no-others --no-cpu-baseline --no-parity > gpurun_out/r2_bench_stage_nt${nt}_t$t.json 2>gpurun_out/r2_bench_stage.err; echo "nt $nt threads $t rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_stage_nt${nt}_t$t.json')); print(l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['e2e']['pageable']['ms_per_step'], l['e2e']['pageable']['vs_pinned'])"; done; done
timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --extra-kinds > gpurun_out/r2_bench_1080p.json 2>gpurun_out/r2_bench_1080p.err; echo "1080p rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_1080p.json')); print(l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['e2e']['pageable'], l['parity'], l.get('other_distributions'))"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 12 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"; grep -v "^==" gpurun_out/r2_launches_1080p.csv | cut -d, -f5,15 | head -14
