#!/bin/bash
# round 2, eighth GPU call (1 GPU): HEAD baseline -- whole GPU test-suite, smoke, default bench line, 1080p line with the
# other distributions, launch list and full ncu capture (with source) of one launch sequence of the 1080p search stage
mkdir -p gpurun_out
nproc > gpurun_out/r2_host.txt; nvidia-smi topo -m >> gpurun_out/r2_host.txt 2>&1; lscpu | head -30 >> gpurun_out/r2_host.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_smoke.log
timeout 300 python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --extra-kinds > gpurun_out/r2_bench_1080p.json 2>gpurun_out/r2_bench_1080p.err; echo "1080p rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_1080p.json')); print(l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['e2e'].get('pageable'), l['parity'], l.get('other_distributions'))"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 12 --csv --log-file gpurun_out/r2_launches_1080p.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"; grep -v "^==" gpurun_out/r2_launches_1080p.csv | cut -d, -f5,15 | head -14
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 6 -f -o gpurun_out/r2_1080p_full python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full 1080p rc=$?"
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_default.err; cut -c1-1200 gpurun_out/r2_bench_default.json
