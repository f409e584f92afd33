#!/bin/bash
# round-2 call A: engine vs reference-on-GPU parity, library calibration, bench line with the reference arms
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_vs_reference.py -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_vs_ref.log 2>&1
echo "== pytest exit $?"; grep -E "^\[|^  [ER]-|passed|failed|Error" gpurun_out/pytest_vs_ref.log | tail -n 40
timeout 600 python scripts/calib_library.py > gpurun_out/calib_library.txt 2> gpurun_out/calib_library.err
echo "== calib exit $?"; cat gpurun_out/calib_library.txt; tail -n 3 gpurun_out/calib_library.err
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_base16.json 2> gpurun_out/bench_base16.err
echo "== bench exit $?"; tail -c 5000 gpurun_out/bench_base16.json; tail -n 5 gpurun_out/bench_base16.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "== bench ref exit $?"; tail -c 2000 gpurun_out/bench_ref.json
