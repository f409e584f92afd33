#!/bin/bash
# compute-sanitizer over the hand-written kernels (ONE tool per gpurun call: B200_PROFILING.md).
#   bash scripts/sanitize.sh memcheck|racecheck|synccheck
# Runs the kernel-level parity tests (tcgen05 GEMM / conv incl. CTA pairs, TMA reduce-add, 512-row tiles; attention incl. the
# streamed softmax; SIMT kernels) and one tiny UFM-Refine forward (__graft_entry__.smoke) under the tool, after the same
# commands passed without it.  The summary lands in gpurun_out/sanitizer_<tool>.txt (copied to profiles/ by hand).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SEL="gemm_bias or gemm_gelu or rows2 or inplace_residual or conv3x3_implicit or conv3x3_tail or test_attention or layernorm or qk_norm or local_refinement or resize_align"
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "$SEL" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -n 5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 600 python __graft_entry__.py --smoke >> gpurun_out/sanitize_plain.log 2>&1 || { echo "plain smoke failed"; exit 1; }
OUT=gpurun_out/sanitizer_$TOOL.txt
echo "# compute-sanitizer --tool $TOOL: kernel parity tests ($SEL) + smoke()" > $OUT
timeout 2400 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 1 \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -x -k "$SEL" > gpurun_out/sanitize_${TOOL}_tests.log 2>&1
echo "tests: exit $?" >> $OUT
grep -E "ERROR SUMMARY|passed|failed|Invalid|Race reported|hazard" gpurun_out/sanitize_${TOOL}_tests.log | tail -n 20 >> $OUT
timeout 1200 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 1 python __graft_entry__.py --smoke > gpurun_out/sanitize_${TOOL}_smoke.log 2>&1
echo "smoke: exit $?" >> $OUT
grep -E "ERROR SUMMARY|smoke:|Invalid|Race reported|hazard" gpurun_out/sanitize_${TOOL}_smoke.log | tail -n 20 >> $OUT
cat $OUT
