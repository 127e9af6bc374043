#!/bin/bash
# HEAD on 2 GPUs: the multi-GPU tests (skipped on a 1-GPU box), the default bench line and the giant line under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "multi or nccl or comm_band or band" > gpurun_out/r2_head_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2_head_pytest_n2.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
timeout 900 $T bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_head_default_n2.json 2> gpurun_out/r2_head_default_n2.err; echo "default n2 rc=$?"; tail -n 2 gpurun_out/r2_head_default_n2.err
timeout 600 $T bench.py --gpus 2 --workload giant --steps 10 --warmup 3 --no-others > gpurun_out/r2_head_giant_n2.json 2> gpurun_out/r2_head_giant_n2.err; echo "giant n2 rc=$?"; tail -n 2 gpurun_out/r2_head_giant_n2.err
python - <<'PY'
import json
for f in ("r2_head_default_n2", "r2_head_giant_n2"):
    try:
        l = json.load(open(f"gpurun_out/{f}.json"))
        print(f, l["ms_per_step"], l["value"], l["e2e"].get("ms_per_step"), l["parity"]["bit_exact"], l["clocks"])
    except Exception as e:
        print(f, "unreadable", e)
PY
