#!/bin/bash
# 512-row tiles with the early release of both accumulators in the bf16 epilogue (packed row block 0 beside raw row block 1,
# setmaxnreg 72 / 216): kernel tests, micro-benchmark of the rows2 modes, in-step A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "gemm or conv or linear" > gpurun_out/pytest_x.log 2>&1
echo "== pytest exit $?"; tail -n 4 gpurun_out/pytest_x.log
timeout 600 python scripts/gemm_bench.py --no-conv --variants 1:3:1:0:0:1,1:3:1:27:0:1,1:3:1:3:0:1,1:3:1:7:0:1 > gpurun_out/gemm_bench_x.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_x.txt
for rep in 1 2; do
for m in 27 3 7 0; do
  UFM_ROWS2=$m timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_x_m${m}_$rep.json 2>gpurun_out/bench_x.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_x_m${m}_$rep.json").read().strip().splitlines()[-1])
print("rows2=$m rep=$rep", round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "attention")})
PY
done
done
