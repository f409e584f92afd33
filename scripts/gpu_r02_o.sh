#!/bin/bash
# concurrent branches for <= 4 pairs: bit-identity test, parity suite, batch sweep with the branches off / on
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -x -k "branches or chunking or golden or api" > gpurun_out/pytest_branches.log 2>&1
echo "== pytest exit $?"; tail -n 12 gpurun_out/pytest_branches.log
for M in 0 1; do
  UFM_BRANCHES=$M timeout 600 python scripts/sweep_batch.py --batches 1,2,4,8 > gpurun_out/sweep_branches$M.csv 2>&1
  echo "== sweep branches=$M exit $?"; cat gpurun_out/sweep_branches$M.csv
done
UFM_BRANCHES=1 timeout 600 python scripts/sweep_batch.py --batches 1,2,4 --host > gpurun_out/sweep_branches1_host.csv 2>&1
cat gpurun_out/sweep_branches1_host.csv
UFM_BRANCHES=1 timeout 600 python scripts/sweep_batch.py --workload refine16 --batches 1,2,4 > gpurun_out/sweep_branches1_refine.csv 2>&1
cat gpurun_out/sweep_branches1_refine.csv
UFM_BRANCHES=0 timeout 600 python scripts/sweep_batch.py --workload refine16 --batches 1,2,4 > gpurun_out/sweep_branches0_refine.csv 2>&1
cat gpurun_out/sweep_branches0_refine.csv
