#!/bin/bash
# Kernel-level parity on the GPU box, one pytest process per kernel family so that a trapped kernel
# (poisoned CUDA context) does not hide the results of the others.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for grp in "gemm" "conv3x3" "attention" "layernorm or qk_norm or resize or shuffle or im2col or refinement"; do
  tag=$(echo "$grp" | cut -d' ' -f1)
  timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "$grp" -p no:cacheprovider > gpurun_out/k_$tag.log 2>&1
  echo "== $tag exit $?" ; tail -n 25 gpurun_out/k_$tag.log
done
