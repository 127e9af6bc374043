#!/bin/bash
# round 2, eleventh GPU call: origin-tile near stage with its defaults (2 warps, 48 registers, 12 rings): tests, the launch-sequence
# size, the default bench line, ncu of the 1080p sequence
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
for sub in 2 4 8 16; do STEPTH_BENCH_SUB=$sub timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_sub$sub.json 2>gpurun_out/r2_sub.err; python -c "
import json; l=json.load(open('gpurun_out/r2_sub$sub.json')); print('SUB $sub', l['ms_per_step'], l['roofline']['launch_ms'])"; done
timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --extra-kinds > gpurun_out/r2_bench_1080p.json 2>gpurun_out/r2_bench_1080p.err; echo "1080p rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_1080p.json')); print(l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['parity']['bit_exact'], l.get('other_distributions'))"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 12 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"; grep -v "^==" gpurun_out/r2_launches_1080p.csv | cut -d, -f5,15 | head -8
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 6 -f -o gpurun_out/r2_otile_full python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"search_|tile_planes|coarse_transpose" -c 4 -f -o gpurun_out/r2_batch4k_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full4k.log 2>&1; echo "ncu full 4k rc=$?"
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_default.err; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_default.json')); print(l['ms_per_step'], l['e2e']['ms_per_step'], l['parity']['bit_exact']); print({k:(v['ms_per_step'], v['e2e']['ms_per_step'] if v['e2e'] else None, v['parity']['bit_exact']) for k,v in l['other_workloads'].items()})"
