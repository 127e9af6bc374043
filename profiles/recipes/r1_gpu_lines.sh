#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --extra-kinds > gpurun_out/bench_1080p.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --workload 4k_fused --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_4k_fused.json 2> gpurun_out/bench4k.err; echo "bench4k rc=$?"
timeout 600 python bench.py --workload batch4k --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_batch4k_n1.json 2> gpurun_out/benchb4k.err; echo "batch4k rc=$?"
