#!/bin/bash
# round 2, fourteenth GPU call: K0 strips through a tensor map (tests, K0 time with / without), register cap 56, chunking of the batch call
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest_gpu.log
B="python bench.py --workload 1080p --steps 10 --warmup 3 --no-others --no-cpu-baseline --no-e2e"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', l['ms_per_step'], l['roofline']['launch_ms'], l['parity']['bit_exact'])" || tail -3 gpurun_out/r2_v_$name.err; }
run tmap1 STEPTH_K0_TMAP=1
run tmap0 STEPTH_K0_TMAP=0
run mb18 STEPTH_LIB=stepth_b200/libstepth_cuda_w2mb18.so
run tmap1b STEPTH_K0_TMAP=1
for k in 0 1; do STEPTH_K0_TMAP=$k timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tile_planes" -c 4 --csv --log-file gpurun_out/r2_k0_tmap$k.csv python bench.py --workload 1080p --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch.log 2>&1; grep -v "^==" gpurun_out/r2_k0_tmap$k.csv | cut -d, -f5,15 | tail -3; done
for ch in 8 16 32; do STEPTH_BATCH_CHUNKS=$ch timeout 600 python bench.py --steps 5 --warmup 3 --no-others --no-cpu-baseline --no-parity > gpurun_out/r2_chunks$ch.json 2>gpurun_out/r2_chunks.err; python -c "
import json; l=json.load(open('gpurun_out/r2_chunks$ch.json')); print('chunks $ch', l['ms_per_step'], l['e2e']['ms_per_step'], l['e2e'].get('frac_of_link_ceiling'), l['e2e']['pageable']['ms_per_step'], l['e2e']['pageable']['vs_pinned'])"; done
for t in 12 16; do STEPTH_STAGE_THREADS=$t timeout 600 python bench.py --steps 5 --warmup 3 --no-others --no-cpu-baseline --no-parity > gpurun_out/r2_st$t.json 2>gpurun_out/r2_st.err; python -c "
import json; l=json.load(open('gpurun_out/r2_st$t.json')); print('stage threads $t', l['e2e']['ms_per_step'], l['e2e']['pageable']['ms_per_step'], l['e2e']['pageable']['vs_pinned'])"; done
