#!/bin/bash
# the round's evidence run: tests, smoke, bench lines, ncu launch list, ncu full capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --extra-kinds > gpurun_out/bench_1080p.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/benchref.err; echo "ref rc=$?"
timeout 600 python bench.py --workload 4k_fused --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_4k_fused.json 2> gpurun_out/bench4k.err; echo "bench4k rc=$?"; tail -3 gpurun_out/bench4k.err
timeout 600 python bench.py --workload batch4k --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_batch4k_n1.json 2> gpurun_out/benchb4k.err; echo "batch4k rc=$?"; tail -3 gpurun_out/benchb4k.err
timeout 600 python bench.py --workload giant --steps 3 --warmup 3 > gpurun_out/bench_giant_n1.json 2> gpurun_out/benchg.err; echo "giant rc=$?"; tail -3 gpurun_out/benchg.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_|mask_apply|depth_split" -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"search_|tile_planes|coarse_transpose|finish_" -c 6 -f -o gpurun_out/hier_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
STEPTH_TRACE=1 timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_trace.json 2> gpurun_out/trace.err; echo "trace rc=$?"; tail -12 gpurun_out/trace.err
timeout 300 python bench.py --pairs 1 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_1080p_one_pair.json 2> gpurun_out/b1.err; echo "one pair rc=$?"
