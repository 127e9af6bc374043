#!/bin/bash
for v in "STEPTH_BATCH_CHUNKS=8" "STEPTH_BATCH_CHUNKS=11" "STEPTH_BATCH_CHUNKS=16"; do
  env $v python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_x.json 2>gpurun_out/e2e_x.err || tail -3 gpurun_out/e2e_x.err
  python -c "
import json; d=json.load(open('gpurun_out/e2e_x.json')); print('$v resident ms', d['ms_per_step'], 'e2e ms', d['e2e']['ms_per_step'], 'e2e value', d['e2e']['value'])"
done
