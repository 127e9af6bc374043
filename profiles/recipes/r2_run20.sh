#!/bin/bash
# round 2, twentieth GPU call: decode / encode on the device (nvJPEG) tests; launch sequences of 32 pairs
mkdir -p gpurun_out
ls /usr/local/cuda/lib64/libnvjpeg* > gpurun_out/r2_nvjpeg.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -k "jpeg or assets" > gpurun_out/r2_pytest_codec.log 2>&1; echo "pytest(codec) rc=$?"; tail -15 gpurun_out/r2_pytest_codec.log
for sub in 16 32; do STEPTH_BENCH_SUB=$sub timeout 300 python bench.py --workload 1080p --pairs 32 --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_sub$sub.json 2>gpurun_out/r2_sub.err; python -c "
import json; l=json.load(open('gpurun_out/r2_sub$sub.json')); print('1080p x32 SUB $sub', l['ms_per_step'], l['roofline']['launch_ms'])"; done
for sub in 16 32; do STEPTH_BENCH_SUB=$sub timeout 300 python bench.py --steps 5 --warmup 3 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_sub$sub.json 2>gpurun_out/r2_sub.err; python -c "
import json; l=json.load(open('gpurun_out/r2_sub$sub.json')); print('batch4k SUB $sub', l['ms_per_step'], l['roofline']['launch_ms'])"; done
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
