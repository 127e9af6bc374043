#!/bin/bash
# last call of the round (a few GPU-minutes left): the library rebuilt in a re-created container -- smoke, the no-alloc test alone,
# a handful of parity tests
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/last_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/last_smoke.log
timeout 100 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/last_noalloc.log 2>&1; echo "noalloc rc=$?"; tail -n 2 gpurun_out/last_noalloc.log
timeout 150 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "assets or blocks or dense_vs_oracle" > gpurun_out/last_parity.log 2>&1; echo "parity rc=$?"; tail -n 2 gpurun_out/last_parity.log
