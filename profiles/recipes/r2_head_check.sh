#!/bin/bash
# HEAD check: the whole GPU test-suite, smoke, the default bench line (short)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.log
timeout 1500 python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_bench_default.json')); print(l['ms_per_step'], l['e2e']['ms_per_step'], l['parity']['bit_exact'], l['gpu_launches'], l['clocks']); print({k:(v['ms_per_step'], v['parity']['bit_exact']) for k,v in l['other_workloads'].items()})"
