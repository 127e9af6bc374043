#!/bin/bash
for v in "STEPTH_NEAR_RINGS=6" "STEPTH_FINISH_STREAMING=0" "STEPTH_TWO_STREAMS=0" "STEPTH_BATCH_CHUNKS=2" "STEPTH_TWO_STREAMS=0 STEPTH_FINISH_STREAMING=0"; do
  env $v python bench.py --workload 4k_fused --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_x.json 2>gpurun_out/e2e_x.err || tail -3 gpurun_out/e2e_x.err
  python -c "
import json; d=json.load(open('gpurun_out/e2e_x.json')); print('$v resident ms', d['ms_per_step'], 'e2e ms', d['e2e']['ms_per_step'], 'e2e value', d['e2e']['value'])"
done
