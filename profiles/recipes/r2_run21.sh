#!/bin/bash
# round 2, twenty-first GPU call: small launches -- the walking near kernel against the origin-tile stage; launch list of the 4K fused tail
mkdir -p gpurun_out
run() { name=$1; wl=$2; pairs=$3; shift; shift; shift; env "$@" timeout 300 python bench.py --workload $wl --pairs $pairs --steps 30 --warmup 5 --no-others --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2_v_$name.json 2>gpurun_out/r2_v_$name.err; python -c "
import json; l=json.load(open('gpurun_out/r2_v_$name.json')); print('$name', '$wl', $pairs, l['ms_per_step'], l['roofline']['launch_ms'])" || tail -3 gpurun_out/r2_v_$name.err; }
for p in 1 2 3 4; do run otile 1080p $p X=1; run walk 1080p $p STEPTH_NEAR_TILE=0; done
run otile 4k_fused 1 X=1; run walk 4k_fused 1 STEPTH_NEAR_TILE=0
run otile 4k_fused 2 X=1; run walk 4k_fused 2 STEPTH_NEAR_TILE=0
run otile readme_dense 1 X=1; run walk readme_dense 1 STEPTH_NEAR_TILE=0
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_|depth_split|mask_apply" -c 16 --csv --log-file gpurun_out/r2_launches_batch4k.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-others --no-parity > gpurun_out/r2_ncu_launch4k.log 2>&1; echo "ncu launches 4k rc=$?"; grep -v "^==" gpurun_out/r2_launches_batch4k.csv | awk -F'","' '{print $5, $NF}' | tr -d '"' | cut -c1-60 | head -18
