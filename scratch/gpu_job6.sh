#!/bin/bash
mkdir -p gpurun_out
M="gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__data_pipe_lsu_wavefronts_mem_lg.sum"
for kind in nat; do
timeout 600 ncu --metrics $M --clock-control none -k regex:"search_|tile_planes|coarse_tr|finish" -c 6 --csv --log-file gpurun_out/k_$kind.csv python bench.py --kind $kind --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/k_$kind.log 2>&1; echo "$kind rc=$?"
done
