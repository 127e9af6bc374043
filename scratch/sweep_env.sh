#!/bin/bash
# usage: sweep_env.sh "VAR=1 VAR2=3" "..." : one bench run per argument (env assignments), prints the search-stage ms
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs python bench.py --steps 4 --warmup 2 --extra-kinds --no-cpu-baseline --no-e2e > gpurun_out/sw_$i.json 2> gpurun_out/sw_$i.err || tail -3 gpurun_out/sw_$i.err
  python - "$envs" gpurun_out/sw_$i.json <<'PY'
import json,sys
d=json.load(open(sys.argv[2])); o=d.get('other_distributions',{})
print(f"{sys.argv[1]:60s} nat {d['roofline']['launch_ms']:8.3f}  geo {o.get('geo',{}).get('search_ms_per_launch')}  worst {o.get('worst',{}).get('search_ms_per_launch')}")
PY
done
