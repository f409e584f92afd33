#!/bin/bash
# in-step A/B of the attention polynomial share under the power cap, interleaved on one box
mkdir -p gpurun_out
for rep in 1 2; do
for m in 132 128 134 136 131; do
  UFM_ATTN_POLY=$m timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_s_m${m}_$rep.json 2>gpurun_out/bench_s.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_s_m${m}_$rep.json").read().strip().splitlines()[-1])
print("poly=$m rep=$rep", round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "attention")})
PY
done
done
