#!/bin/bash
# rows2 with the remainder folded into the same launch (half items) + packed GELU: kernel tests, micro-benchmark, step
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "gemm or conv or linear" > gpurun_out/pytest_p.log 2>&1
echo "== pytest exit $?"; tail -n 15 gpurun_out/pytest_p.log
timeout 600 python scripts/gemm_bench.py --no-conv --variants 1:1:1:0:0:0,1:1:1:3:0:0,1:1:1:11:0:0,1:1:1:7:0:0,1:1:1:0:0:1 > gpurun_out/gemm_bench_p.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_p.txt
for cfg in "0 0" "11 0" "0 1" "11 1" "3 1" "0 0"; do
  set -- $cfg
  UFM_ROWS2=$1 UFM_GELU_PACKED=$2 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_p_r$1_g$2.json 2>gpurun_out/bench_p_r$1_g$2.err
  echo "== bench rows2=$1 gelu_packed=$2 exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_p_r$1_g$2.json").read().strip().splitlines()[-1])
    kf = d["kernel_families"]
    print(d["value"], d["ms_per_step"], d["clocks"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in kf.items()})
except Exception as e:
    print("parse failed", e)
PY
done
