#!/bin/bash
# new defaults (rows2 = 11 with folded half items, packed GELU): whole GPU suite, micro-benchmark, default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_q.log 2>&1
echo "== pytest exit $?"; tail -n 6 gpurun_out/pytest_q.log
timeout 600 python scripts/gemm_bench.py --no-conv --variants 1:1:1:0:0:0,1:1:1:11:0:1,1:1:1:15:0:1,1:1:1:3:0:1 > gpurun_out/gemm_bench_q.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_q.txt
for i in 1 2; do
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_q$i.json 2>gpurun_out/bench_q$i.err
echo "== bench exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_q$i.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["ms_per_step"], d["clocks"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items()})
PY
done
