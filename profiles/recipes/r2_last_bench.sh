#!/bin/bash
# the driver's own command (no flags) at HEAD, timed by the wall clock: does the default line still finish within minutes?
mkdir -p gpurun_out
s=$(date +%s)
timeout 230 python bench.py > gpurun_out/last_bench_default.json 2> gpurun_out/last_bench_default.err; echo "bench rc=$? wall=$(( $(date +%s) - s )) s"
tail -n 2 gpurun_out/last_bench_default.err
python -c "
import json; l=json.load(open('gpurun_out/last_bench_default.json')); print(l['ms_per_step'], l['value'], l['e2e']['ms_per_step'], l['e2e'].get('frac_of_link_ceiling'), l['parity']['bit_exact'], l['gpu_launches'], l['clocks']); print({k:(v['ms_per_step'], v['parity']['bit_exact']) for k,v in l['other_workloads'].items()})"
