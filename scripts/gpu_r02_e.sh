#!/bin/bash
# round-2 call E: rows2 GEMM with the early-release epilogue, attention with delayed P stores
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "gemm or attention or conv" > gpurun_out/pytest_kernels.log 2>&1
echo "== pytest kernels exit $?"; tail -n 5 gpurun_out/pytest_kernels.log
timeout 600 python scripts/gemm_bench.py --variants 1:1:1:0:0,1:1:1:3:0,1:1:1:7:0,1:1:1:7:4 --reps 20 > gpurun_out/gemm_bench_rows2.txt 2>&1
echo "== gemm_bench exit $?"; cat gpurun_out/gemm_bench_rows2.txt
timeout 300 python scripts/attn_bench.py 4,132 > gpurun_out/attn_bench.txt 2>&1
echo "== attn_bench exit $?"; cat gpurun_out/attn_bench.txt | tail -n 8
for r in 0 3 7; do
UFM_ROWS2=$r timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_base16_r$r.json 2> gpurun_out/bench_base16_r$r.err
echo "== bench rows2=$r exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_base16_r$r.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])
for k,v in d['kernel_families'].items(): print(k, round(v['ms_per_step'],3), round(v['achieved'],1), round(v['frac'],3))
PY
done
