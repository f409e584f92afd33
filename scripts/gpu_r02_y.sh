#!/bin/bash
# table-driven uint8 patch gather + 32-bit index copies: parity / kernel tests, move family in the step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_y.log 2>&1
echo "== pytest exit $?"; tail -n 5 gpurun_out/pytest_y.log
for rep in 1 2; do
for g in 1 0; do
  if [ $g = 1 ]; then export UFM_PATCH_GENERIC=1; else unset UFM_PATCH_GENERIC; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_y_g${g}_$rep.json 2>gpurun_out/bench_y.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_y_g${g}_$rep.json").read().strip().splitlines()[-1])
print("generic_patch=$g rep=$rep", round(d["value"], 2), round(d["e2e"]["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 3), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "move")})
PY
done
done
unset UFM_PATCH_GENERIC
timeout 300 python scripts/sweep_batch.py --batches 1,2,4 > gpurun_out/sweep_y.csv 2>&1; cat gpurun_out/sweep_y.csv
