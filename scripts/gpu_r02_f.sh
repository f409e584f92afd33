#!/bin/bash
# round-2 call F: full GPU suite, bench lines (base16 default incl. reference arms, foreign16), small-batch sweep (graph replay)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x -s > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench exit $?"; tail -c 1500 gpurun_out/bench_base16.json; tail -n 3 gpurun_out/bench_base16.err
timeout 600 python bench.py --steps 10 --warmup 3 --workload foreign16 --no-cpu --no-torch-gpu > gpurun_out/bench_foreign16.json 2> gpurun_out/bench_foreign16.err
echo "== bench foreign exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_foreign16.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'])
for k,v in d['kernel_families'].items(): print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
tail -n 3 gpurun_out/bench_foreign16.err
timeout 600 python scripts/sweep_batch.py --batches 1,2,4,8,16,32 > gpurun_out/sweep_batch_device.csv 2>&1
echo "== sweep exit $?"; cat gpurun_out/sweep_batch_device.csv
UFM_NO_GRAPH=1 timeout 600 python scripts/sweep_batch.py --batches 1,2,4 > gpurun_out/sweep_batch_device_nograph.csv 2>&1
echo "== sweep nograph exit $?"; cat gpurun_out/sweep_batch_device_nograph.csv
timeout 600 python scripts/sweep_batch.py --batches 1,2,4,16 --host > gpurun_out/sweep_batch_host.csv 2>&1
echo "== sweep host exit $?"; cat gpurun_out/sweep_batch_host.csv
