#!/bin/bash
# usage: sweep.sh "v1 v2 ..."  -> bench nat/geo/worst per STEPTH_EXPORT_BELOW value
for v in $1; do
  STEPTH_EXPORT_BELOW=$v python bench.py --steps 3 --warmup 3 --extra-kinds --no-cpu-baseline --no-e2e > gpurun_out/sw_$v.json 2> gpurun_out/sw_$v.err || tail -3 gpurun_out/sw_$v.err
  python -c "
import json; d=json.load(open('gpurun_out/sw_$v.json')); o=d.get('other_distributions',{}); print('export_below=$v nat ms', d['ms_per_step'], 'geo', o.get('geo',{}).get('ms_per_step'), 'worst', o.get('worst',{}).get('ms_per_step'))"
done
