#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload giant --steps 3 --warmup 3 > gpurun_out/giant_n1.json 2> gpurun_out/giant_n1.err; echo "n1 rc=$?"; cat gpurun_out/giant_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload giant --steps 3 --warmup 3 > gpurun_out/giant_n2.json 2> gpurun_out/giant_n2.err; echo "n2 rc=$?"; cat gpurun_out/giant_n2.json; tail -3 gpurun_out/giant_n2.err
