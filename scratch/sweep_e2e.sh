#!/bin/bash
for c in 2 4 8 16; do
  STEPTH_BATCH_CHUNKS=$c python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_$c.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/e2e_$c.json')); print('chunks $c resident ms', d['ms_per_step'], 'e2e ms', d['e2e']['ms_per_step'], 'e2e value', d['e2e']['value'])"
done
