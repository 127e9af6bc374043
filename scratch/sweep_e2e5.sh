#!/bin/bash
for wl in 1080p 4k_fused batch4k; do
  python bench.py --workload $wl --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_x.json 2>gpurun_out/e2e_x.err || tail -3 gpurun_out/e2e_x.err
  python -c "
import json; d=json.load(open('gpurun_out/e2e_x.json')); print('$wl resident ms', d['ms_per_step'], 'e2e ms', d['e2e']['ms_per_step'], 'e2e value', d['e2e']['value'])"
done
