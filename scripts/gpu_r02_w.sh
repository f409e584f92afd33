#!/bin/bash
# 2-GPU check of the final binary: NCCL sharded-predict test, weak-scaling bench line under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dp_nccl.py -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_w.log 2>&1
echo "== pytest dp_nccl exit $?"; tail -n 4 gpurun_out/pytest_w.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_w_2gpu.json 2>gpurun_out/bench_w_2gpu.err
echo "== bench 2 gpus exit $?"; tail -n 1 gpurun_out/bench_w_2gpu.json | cut -c1-900
