#!/bin/bash
# One GPU-box pass: parity tests, bench line, then (only after the plain command exited 0) the ncu launch list of one
# forward pass and one --set full capture of the top kernels.  Logs land in gpurun_out/.
mkdir -p gpurun_out
WL=${1:-base16}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
if [ -z "$SKIP_TESTS" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_gpu.log 2>&1
  echo "== pytest exit $?"; tail -n 8 gpurun_out/pytest_gpu.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 --workload $WL > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err
echo "== bench exit $?"; tail -c 4000 gpurun_out/bench_$WL.json; tail -n 5 gpurun_out/bench_$WL.err
if [ -n "$SKIP_NCU" ]; then exit 0; fi
timeout 600 python scripts/one_step.py --workload $WL > gpurun_out/plain_$WL.log 2>&1
rc=$?
echo "== one_step exit $rc"; tail -n 3 gpurun_out/plain_$WL.log
[ $rc -ne 0 ] && exit 0
L=$(grep -o 'launches_per_pass [0-9]*' gpurun_out/plain_$WL.log | awk '{print $2}')
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((2 * L)) -c $L --csv \
  --log-file gpurun_out/launches_$WL.csv python scripts/one_step.py --workload $WL > gpurun_out/ncu_list_$WL.log 2>&1
echo "== ncu list exit $?"; tail -n 3 gpurun_out/ncu_list_$WL.log
# --set full: the four encoder-block GEMMs + attention of the third pass (skip counts kernels matching -k only)
NK=${NCU_FULL_COUNT:-12}
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"${NCU_FULL_REGEX:-gemm_kernel|attn_kernel}" \
  -s ${NCU_FULL_SKIP:-0} -c $NK -o gpurun_out/prof_$WL python scripts/one_step.py --workload $WL --warm 0 ${NCU_FULL_ARGS} \
  > gpurun_out/ncu_full_$WL.log 2>&1
echo "== ncu full exit $?"; tail -n 3 gpurun_out/ncu_full_$WL.log
ls -la gpurun_out
