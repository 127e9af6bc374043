#!/bin/bash
# round 2, twelfth GPU call: device discretiser (tests + readme_blocks line), warp granularity of the origin-tile stage on small launches
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest_gpu.log
for wl in readme_blocks readme_dense; do timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-others --no-cpu-baseline > gpurun_out/r2_$wl.json 2>gpurun_out/r2_$wl.err; echo "$wl rc=$?"; python -c "
import json; l=json.load(open('gpurun_out/r2_$wl.json')); print('$wl', l['ms_per_step'], l['roofline']['launch_ms'], l['e2e']['ms_per_step'], l['e2e'].get('pageable',{}).get('ms_per_step'), l['parity'])" || tail -5 gpurun_out/r2_$wl.err; done
for nit in 8 4 2 1; do
STEPTH_OT_NIT=$nit timeout 300 python bench.py --workload readme_dense --steps 20 --warmup 3 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_nit.json 2>gpurun_out/r2_nit.err; python -c "
import json; l=json.load(open('gpurun_out/r2_nit.json')); print('readme_dense nit $nit', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])"
STEPTH_OT_NIT=$nit timeout 300 python bench.py --workload 1080p --pairs 1 --steps 30 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_nit.json 2>gpurun_out/r2_nit.err; python -c "
import json; l=json.load(open('gpurun_out/r2_nit.json')); print('1080p one pair nit $nit', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])"
STEPTH_OT_NIT=$nit timeout 300 python bench.py --workload 4k_fused --pairs 1 --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_nit.json 2>gpurun_out/r2_nit.err; python -c "
import json; l=json.load(open('gpurun_out/r2_nit.json')); print('4k one pair nit $nit', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])"
done
STEPTH_NEAR_TILE=0 timeout 300 python bench.py --workload readme_dense --steps 20 --warmup 3 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_nit.json 2>gpurun_out/r2_nit.err; python -c "
import json; l=json.load(open('gpurun_out/r2_nit.json')); print('readme_dense walk', l['ms_per_step'], l['roofline']['launch_ms'])"
STEPTH_NEAR_TILE=0 timeout 300 python bench.py --workload 4k_fused --pairs 1 --steps 20 --warmup 5 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_nit.json 2>gpurun_out/r2_nit.err; python -c "
import json; l=json.load(open('gpurun_out/r2_nit.json')); print('4k one pair walk', l['ms_per_step'], l['roofline']['launch_ms'])"
