#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err; echo "bench rc=$?"; tail -3 gpurun_out/b.err
python -c "
import json; d=json.load(open('gpurun_out/b.json')); print('nat step', d['ms_per_step'], 'search', d['roofline']['launch_ms'])"
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:"finish_" -c 6 --csv --log-file gpurun_out/fin.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_fin.log 2>&1; echo "ncu rc=$?"
grep -v "^==" gpurun_out/fin.csv | cut -d, -f5,13- | tail -18
