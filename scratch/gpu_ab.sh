#!/bin/bash
for lib in scratch/lib_base.so stepth_b200/libstepth_cuda.so; do
for k in nat geo worst; do
STEPTH_LIB=$lib timeout 600 python bench.py --workload 4k_fused --kind $k --steps 4 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/b4k.json 2> gpurun_out/b4k.err; tail -3 gpurun_out/b4k.err
python - $k $lib <<'PY'
import json,sys
d=json.load(open('gpurun_out/b4k.json')); print(sys.argv[2], "4k_fused", sys.argv[1], "step ms", d['ms_per_step'], "search", d['roofline']['launch_ms'])
PY
done; done
