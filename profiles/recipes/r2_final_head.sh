#!/bin/bash
# the round's last GPU call: the whole GPU suite (with the graph-capture test) and the driver's own bench command at HEAD
mkdir -p gpurun_out
timeout 100 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/final_pytest_gpu.log
s=$(date +%s)
timeout 110 python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err; echo "bench rc=$? wall=$(( $(date +%s) - s )) s"
tail -n 2 gpurun_out/final_bench_default.err
python -c "
import json; l=json.load(open('gpurun_out/final_bench_default.json')); print(l['ms_per_step'], l['value'], l['e2e']['ms_per_step'], l['e2e'].get('frac_of_link_ceiling'), l['parity']['bit_exact'], l['gpu_launches'], l['clocks']); print({k:(v['ms_per_step'], v['parity']['bit_exact'], v.get('graph_replay',{}).get('ms_per_step')) for k,v in l['other_workloads'].items()})"
