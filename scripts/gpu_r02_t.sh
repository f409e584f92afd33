#!/bin/bash
# L2 prefetch of the x slabs of an in-place reduce-add epilogue: kernel tests, micro-benchmark, in-step A/B
mkdir -p gpurun_out
UFM_RED_PREFETCH=1 timeout 1200 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "gemm or conv or linear" > gpurun_out/pytest_t.log 2>&1
echo "== pytest exit $?"; tail -n 4 gpurun_out/pytest_t.log
timeout 600 python scripts/gemm_bench.py --no-conv --only proj --variants 1:1:1:0:0:1,1:1:1:0:0:3,1:1:1:27:0:1,1:1:1:27:0:3 > gpurun_out/gemm_bench_t.txt 2>&1
timeout 600 python scripts/gemm_bench.py --no-conv --only fc2 --variants 1:1:1:0:0:1,1:1:1:0:0:3,1:1:1:27:0:1,1:1:1:27:0:3 >> gpurun_out/gemm_bench_t.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_t.txt
for rep in 1 2 3; do
for m in 0 1; do
  UFM_RED_PREFETCH=$m timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_t_m${m}_$rep.json 2>gpurun_out/bench_t.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_t_m${m}_$rep.json").read().strip().splitlines()[-1])
print("red_prefetch=$m rep=$rep", round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "attention")})
PY
done
done
