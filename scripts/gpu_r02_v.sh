#!/bin/bash
# CTA pairs for the BN = 128 tiles (UFM_PAIR=3): whole GPU suite under it, in-step A/B on all workloads
mkdir -p gpurun_out
UFM_PAIR=3 timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_v.log 2>&1
echo "== pytest (UFM_PAIR=3) exit $?"; tail -n 6 gpurun_out/pytest_v.log
for rep in 1 2; do
for m in 1 3; do
  UFM_PAIR=$m timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_v_m${m}_$rep.json 2>gpurun_out/bench_v.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_v_m${m}_$rep.json").read().strip().splitlines()[-1])
print("base16 pair=$m rep=$rep", round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "conv3x3", "attention")})
PY
done
done
for wl in refine16 refine16_unet hires4; do
for m in 1 3; do
  UFM_PAIR=$m timeout 600 python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu --no-torch-gpu > gpurun_out/bench_v_${wl}_m${m}.json 2>gpurun_out/bench_v.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/bench_v_${wl}_m${m}.json").read().strip().splitlines()[-1])
print("$wl pair=$m", round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"], {k: (round(v["ms_per_step"], 2), round(v["frac"], 3)) for k, v in d["kernel_families"].items() if k in ("gemm", "conv3x3", "attention")})
PY
done
done
