#!/bin/bash
# round-2 call G: parity tests of the new configurations, then compute-sanitizer memcheck (one tool per call)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -x -k "reg or hooks or randidx or golden or chunking" > gpurun_out/pytest_newcfg.log 2>&1
echo "== pytest new configs exit $?"; tail -n 4 gpurun_out/pytest_newcfg.log
bash scripts/sanitize.sh memcheck
