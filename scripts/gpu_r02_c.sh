#!/bin/bash
# round-2 call C: rows2 GEMM tiles (tests + micro-benchmark), warp-role shift, attention source-level profile
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "gemm" > gpurun_out/pytest_gemm.log 2>&1
echo "== pytest gemm exit $?"; tail -n 5 gpurun_out/pytest_gemm.log
timeout 600 python scripts/gemm_bench.py --variants 1:1:1:0:0,1:1:1:1:0,1:1:1:3:0,1:1:1:7:0,1:1:1:0:4,1:1:1:3:4 --reps 20 > gpurun_out/gemm_bench_rows2.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_rows2.txt
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_base16.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])
for k,v in d['kernel_families'].items(): print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
timeout 300 python scripts/attn_bench.py 132 > gpurun_out/attn_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_kernel -s 26 -c 1 -o gpurun_out/attn_v132 python scripts/attn_bench.py 132 > gpurun_out/attn_ncu.log 2>&1
echo "== ncu attn exit $?"; tail -n 2 gpurun_out/attn_ncu.log
