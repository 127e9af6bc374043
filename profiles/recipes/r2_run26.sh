#!/bin/bash
# round 2, twenty-sixth GPU call: the whole GPU test-suite through a -DSTEPTH_DEBUG_CHECKS build (index checks of the search kernels;
# compute-sanitizer is closed on this pool), the release build's 1080p line beside it
mkdir -p gpurun_out
STEPTH_LIB=stepth_b200/libstepth_cuda_dbg.so timeout 1500 python -m pytest tests -m gpu -q -k "not cpp_host" > gpurun_out/r2_pytest_dbg.log 2>&1; echo "pytest(debug checks) rc=$?"; tail -3 gpurun_out/r2_pytest_dbg.log
for wl in 1080p 4k_fused giant; do STEPTH_LIB=stepth_b200/libstepth_cuda_dbg.so timeout 600 python bench.py --workload $wl --steps 3 --warmup 1 --no-others --no-cpu-baseline > gpurun_out/r2_dbg_$wl.json 2>gpurun_out/r2_dbg.err; echo "$wl (debug checks) rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_dbg_$wl.json')); print(l['parity']['bit_exact'])"; done
timeout 300 python bench.py --workload 1080p --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_v_rel.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/r2_v_rel.json')); print('release 1080p', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])"
