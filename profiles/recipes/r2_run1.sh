#!/bin/bash
# round 2, first GPU call: the GPU test-suite with the new full-size parity cases, smoke, the default bench line
mkdir -p gpurun_out
nproc > gpurun_out/r2_host.txt; nvidia-smi topo -m >> gpurun_out/r2_host.txt 2>&1; numactl -H >> gpurun_out/r2_host.txt 2>&1; lscpu | head -30 >> gpurun_out/r2_host.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_default.err; cut -c1-1500 gpurun_out/r2_bench_default.json
