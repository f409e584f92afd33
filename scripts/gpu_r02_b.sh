#!/bin/bash
# round-2 call B: full GPU test suite with the streamed attention step, attention variants, cuBLAS / own GEMM ncu probe
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python scripts/attn_bench.py 4,132,131,130,128 > gpurun_out/attn_bench.txt 2>&1
echo "== attn_bench exit $?"; cat gpurun_out/attn_bench.txt | tail -n 24
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x -s > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_base16.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])
for k,v in d['kernel_families'].items(): print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
timeout 300 python scripts/gemm_probe.py > gpurun_out/gemm_probe_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:"gemm|nvjet|cutlass|sm100|xmma" -c 12 -o gpurun_out/gemm_probe python scripts/gemm_probe.py > gpurun_out/gemm_probe_ncu.log 2>&1
echo "== ncu gemm probe exit $?"; tail -n 3 gpurun_out/gemm_probe_ncu.log
ls -la gpurun_out | head -40
