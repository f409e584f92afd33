#!/bin/bash
# round-2 call D: source-level ncu profile of the rows2 GEMM tile (enc.qkv / enc.fc2 shapes) next to the 256-row kernel
mkdir -p gpurun_out
timeout 300 python scripts/gemm_probe.py --shapes enc.qkv,enc.fc2 > gpurun_out/gemm_probe_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_kernel" -c 4 -o gpurun_out/gemm_rows2 python scripts/gemm_probe.py --shapes enc.qkv,enc.fc2 > gpurun_out/gemm_rows2_ncu.log 2>&1
echo "== ncu exit $?"; tail -n 3 gpurun_out/gemm_rows2_ncu.log
