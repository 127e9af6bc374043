#!/bin/bash
# tests + default bench + the other workloads (N=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --extra-kinds > gpurun_out/bench_1080p.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
cat gpurun_out/bench_1080p.json
timeout 600 python bench.py --workload 4k_fused --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_4k_fused.json 2> gpurun_out/bench4k.err; echo "bench4k rc=$?"; tail -3 gpurun_out/bench4k.err
cat gpurun_out/bench_4k_fused.json
timeout 600 python bench.py --workload giant --steps 2 --warmup 3 > gpurun_out/bench_giant_n1.json 2> gpurun_out/benchg.err; echo "giant rc=$?"; tail -3 gpurun_out/benchg.err
cat gpurun_out/bench_giant_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_kernel|mask_apply|depth_split" -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
