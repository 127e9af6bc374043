#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"search_|tile_planes|coarse_transpose|finish_|mask_apply|depth_split|histogram|paint" -c 18 --csv --log-file gpurun_out/launches4k.csv python bench.py --workload 4k_fused --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch4k.log 2>&1; echo "rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches4k.csv')) if len(r)>10 and r[0].isdigit()]
acc={}
for r in rows: acc.setdefault((r[0], r[4][:44]), {})[r[-3]] = r[-1]
for (i,k),v in acc.items(): print(i, k, v)
PY
