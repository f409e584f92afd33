#!/bin/bash
# round-2 final evidence pass (one GPU): full GPU suite, all bench workloads, small-batch sweep, ncu launch list + --set full
# captures of the final binary.  Each ncu command runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench base16 exit $?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "== bench reference exit $?"
for WL in refine16 hires4 refine16_unet foreign16; do
  timeout 600 python bench.py --steps 10 --warmup 3 --workload $WL --no-cpu --no-torch-gpu > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err
  echo "== bench $WL exit $?"
done
python - <<'PY'
import json
for wl in ("base16","refine16","hires4","refine16_unet","foreign16"):
    try:
        d=json.loads(open(f'gpurun_out/bench_{wl}.json').read().strip().splitlines()[-1])
        print(wl, round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), d['clocks'].get('sm_mhz'), {k:(round(v['ms_per_step'],2),round(v['frac'],3)) for k,v in d['kernel_families'].items()})
    except Exception as e: print(wl, 'ERR', e)
PY
timeout 600 python scripts/sweep_batch.py --batches 1,2,4,8,16,32 > gpurun_out/sweep_batch_device.csv 2>&1
echo "== sweep exit $?"; cat gpurun_out/sweep_batch_device.csv
# ---- ncu: launch list + --set full of the block kernels (base16), then the SIMT kernels (refine16)
timeout 600 python scripts/one_step.py --workload base16 > gpurun_out/plain_base16.log 2>&1
rc=$?; echo "== one_step exit $rc"; [ $rc -ne 0 ] && exit 0
L=$(grep -o 'launches_per_pass [0-9]*' gpurun_out/plain_base16.log | awk '{print $2}')
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((2 * L)) -c $L --csv \
  --log-file gpurun_out/launches_base16.csv python scripts/one_step.py --workload base16 > gpurun_out/ncu_list_base16.log 2>&1
echo "== ncu list exit $?"
timeout 600 python scripts/one_step.py --workload base16 --warm 0 > gpurun_out/plain_base16_w0.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"gemm_kernel|attn_kernel" -s 0 -c 12 \
  -o gpurun_out/prof_base16 python scripts/one_step.py --workload base16 --warm 0 > gpurun_out/ncu_full_base16.log 2>&1
echo "== ncu full exit $?"
timeout 600 python scripts/one_step.py --workload refine16 --warm 0 > gpurun_out/plain_refine16_w0.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"layernorm_kernel|resize_ac_up_kernel|refine_kernel|desc_shuffle" -s 4 -c 6 \
  -o gpurun_out/prof_simt python scripts/one_step.py --workload refine16 --warm 0 > gpurun_out/ncu_full_simt.log 2>&1
echo "== ncu simt exit $?"
ls -la gpurun_out | head -50
