#!/bin/bash
# 8-GPU weak-scaling bench line of the final binary
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_8gpu.json 2>gpurun_out/bench_8gpu.err
echo "== bench 8 gpus exit $?"; tail -n 1 gpurun_out/bench_8gpu.json | cut -c1-400
