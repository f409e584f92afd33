#!/bin/bash
# round-2 call H (8 GPUs): 2-GPU NCCL test of dp.predict_sharded with the real engine, then the B x G sweep (configs[4])
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi8.txt 2>&1
CUDA_VISIBLE_DEVICES=0,1 timeout 600 python -m pytest tests/test_dp_nccl.py -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_dp_nccl.log 2>&1
echo "== pytest dp nccl exit $?"; tail -n 3 gpurun_out/pytest_dp_nccl.log
: > gpurun_out/sweep_dp.csv; [ -f profiles/r02_sweep_dp_g1.csv ] && cat profiles/r02_sweep_dp_g1.csv >> gpurun_out/sweep_dp.csv
for G in ${GS:-1 2 4 8}; do
  if [ $G -eq 1 ]; then
    timeout 900 python scripts/sweep_dp.py >> gpurun_out/sweep_dp.csv 2> gpurun_out/sweep_dp_g$G.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29500 + G)) \
      scripts/sweep_dp.py >> gpurun_out/sweep_dp.csv 2> gpurun_out/sweep_dp_g$G.err
  fi
  echo "== sweep G=$G exit $?"; tail -n 2 gpurun_out/sweep_dp_g$G.err
done
cat gpurun_out/sweep_dp.csv
