#!/bin/bash
# the README pair's resident step captured into a CUDA graph by the caller and replayed (bench.py: graph_replay)
mkdir -p gpurun_out
for wl in readme_dense readme_blocks; do
timeout 90 python bench.py --workload $wl --no-others --no-cpu-baseline > gpurun_out/r2_graph_$wl.json 2> gpurun_out/r2_graph_$wl.err; echo "$wl rc=$?"; tail -n 2 gpurun_out/r2_graph_$wl.err
python -c "
import json; l=json.load(open('gpurun_out/r2_graph_$wl.json')); print('$wl', l['ms_per_step'], l.get('graph_replay'), l['parity']['bit_exact'])"
done
timeout 60 python bench.py --workload 1080p --pairs 1 --steps 20 --no-others --no-cpu-baseline --no-e2e > gpurun_out/r2_graph_1080p_one.json 2> gpurun_out/r2_graph_1080p_one.err; echo "1080p one pair rc=$?"
python -c "
import json; l=json.load(open('gpurun_out/r2_graph_1080p_one.json')); print('1080p x1', l['ms_per_step'], l.get('graph_replay'), l['parity']['bit_exact'])"
