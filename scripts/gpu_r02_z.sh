#!/bin/bash
# foreign / different source-target sizes against the reference golden, CLI against the oracle
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_cli.py -m gpu -q -p no:cacheprovider -x -s -k "foreign or cli" > gpurun_out/pytest_z.log 2>&1
echo "== pytest exit $?"; grep -E "shapes|passed|failed|Error|assert" gpurun_out/pytest_z.log | tail -20
