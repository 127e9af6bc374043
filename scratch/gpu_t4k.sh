#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload 4k_fused --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b4k.json 2> gpurun_out/b4k.err; echo "rc=$?"; tail -3 gpurun_out/b4k.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/b4k.json')); print("4k_fused step ms", d['ms_per_step'], "search", d['roofline']['launch_ms'], "e2e ms", d['e2e']['ms_per_step'], "value", d['value'], "e2e", d['e2e']['value'])
PY
