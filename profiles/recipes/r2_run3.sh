#!/bin/bash
# round 2, third GPU call: staging v2 (async copy jobs), RGB entry, new tests; bench lines for stage-thread counts
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "pageable or rgb_main or apply_needs or without_reachable or probe_counter or knobs or batch_and" > gpurun_out/r2_pytest_new.log 2>&1; echo "pytest(new) rc=$?"; tail -5 gpurun_out/r2_pytest_new.log
for t in 4 8 12; do STEPTH_STAGE_THREADS=$t timeout 300 python bench.py --steps 4 --warmup 3 --no-others --no-cpu-baseline --no-parity > gpurun_out/r2_bench_stage_t$t.json 2>gpurun_out/r2_bench_stage_t$t.err; echo "stage threads $t rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_stage_t$t.json')); print(l['e2e']['ms_per_step'], l['e2e'].get('link_floor_ms'), l['e2e']['pageable'])"; done
STEPTH_STAGE_THREADS=8 timeout 300 python bench.py --workload 1080p --steps 6 --warmup 3 --no-others --no-cpu-baseline > gpurun_out/r2_bench_1080p_t8.json 2>gpurun_out/r2_bench_1080p_t8.err; echo "1080p rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_1080p_t8.json')); print(l['ms_per_step'], l['e2e'], l['parity'])"
