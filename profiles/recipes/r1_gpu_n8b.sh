#!/bin/bash
mkdir -p gpurun_out
N=8
timeout 900 python -m pytest tests -m gpu -x -q -k "bands or nccl" > gpurun_out/pytest_gpu_n8.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_n8.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$2 bench.py --gpus $N $3 > gpurun_out/$1_n$N.json 2> gpurun_out/$1_n$N.err; echo "$1 rc=$?"; python - gpurun_out/$1_n$N.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1]); print(d["config"]["workload"], "n", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"], d.get("stage_ms_per_rank", "")[:2] if d.get("stage_ms_per_rank") else "")
PY
}
run bench_1080p 1 "--steps 10 --warmup 3"
run bench_batch4k 2 "--workload batch4k --steps 3 --warmup 3"
run bench_giant 3 "--workload giant --steps 5 --warmup 3"
