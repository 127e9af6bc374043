#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
for k in nat geo worst; do
timeout 600 python bench.py --workload 4k_fused --kind $k --steps 4 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/b4k.json 2> gpurun_out/b4k.err; tail -3 gpurun_out/b4k.err
python - $k <<'PY'
import json,sys
d=json.load(open('gpurun_out/b4k.json')); print("4k_fused", sys.argv[1], "step ms", d['ms_per_step'], "search", d['roofline']['launch_ms'])
PY
done
timeout 600 python bench.py --steps 5 --warmup 3 --extra-kinds --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err; tail -3 gpurun_out/b.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/b.json')); o=d.get('other_distributions',{})
print("1080p nat step", d['ms_per_step'], "search", d['roofline']['launch_ms'], " geo", o['geo']['search_ms_per_launch'], " worst", o['worst']['search_ms_per_launch'])
PY
