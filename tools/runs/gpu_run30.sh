#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin.py -m gpu -q -x --timeout=300 -p no:cacheprovider -k "csr2csc or coo or unit_test or own" > gpurun_out/pytest_f3.log 2>&1
echo "pytest f3 exit $?"; tail -15 gpurun_out/pytest_f3.log
echo finished
