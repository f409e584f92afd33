#!/bin/bash
# small-batch regime: ncu launch list of one pass at 1 and 4 pairs (per-launch durations; plain run first)
mkdir -p gpurun_out
for P in 1 4; do
  timeout 600 python scripts/one_step.py --workload base16 --pairs $P > gpurun_out/plain_b$P.log 2>&1
  rc=$?; echo "== one_step pairs $P exit $rc"; [ $rc -ne 0 ] && exit 0
  L=$(grep -o 'launches_per_pass [0-9]*' gpurun_out/plain_b$P.log | awk '{print $2}')
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s $((2 * L)) -c $L --csv \
    --log-file gpurun_out/launches_b$P.csv python scripts/one_step.py --workload base16 --pairs $P > gpurun_out/ncu_list_b$P.log 2>&1
  echo "== ncu list pairs $P exit $?"
done
