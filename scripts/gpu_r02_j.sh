#!/bin/bash
# round-2 call J: prep_resize with the normalisation table (parity + GB/s), ncu --set full of the resize / refinement kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -x -k "prep_resize or unmap or contract or cli or multiple" > gpurun_out/pytest_prepost.log 2>&1
echo "== pytest prepost exit $?"; tail -n 3 gpurun_out/pytest_prepost.log
timeout 600 python bench.py --steps 10 --warmup 3 --workload foreign16 --no-cpu --no-torch-gpu > gpurun_out/bench_foreign16.json 2> gpurun_out/bench_foreign16.err
echo "== bench foreign exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_foreign16.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'])
for k in ('prep_resize','unmap'): v=d['kernel_families'][k]; print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
timeout 600 python scripts/one_step.py --workload refine16 --warm 0 > gpurun_out/plain_refine16_w0.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"resize_ac_up_kernel|refine_kernel|desc_shuffle" -c 8 \
  -o gpurun_out/prof_simt2 python scripts/one_step.py --workload refine16 --warm 0 > gpurun_out/ncu_full_simt2.log 2>&1
echo "== ncu simt2 exit $?"; tail -n 2 gpurun_out/ncu_full_simt2.log
