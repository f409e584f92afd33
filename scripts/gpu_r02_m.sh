#!/bin/bash
# allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (the new file is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1
echo "== no-alloc alone exit $?"; tail -n 15 gpurun_out/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "== pytest exit $?"; tail -n 15 gpurun_out/pytest_gpu.log
