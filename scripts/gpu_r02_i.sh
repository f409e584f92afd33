#!/bin/bash
# round-2 call I: early PDL trigger: correctness (kernel + parity subset), small-batch sweep, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 4 gpurun_out/pytest_gpu.log
timeout 600 python scripts/sweep_batch.py --batches 1,2,4,8,16 > gpurun_out/sweep_batch_device.csv 2>&1
echo "== sweep exit $?"; cat gpurun_out/sweep_batch_device.csv
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_base16.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])
for k,v in d['kernel_families'].items(): print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
