#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=120 -p no:cacheprovider -k "rowblock" > gpurun_out/pytest_rb.log 2>&1
rc=$?; echo "pytest rb exit $rc"; tail -12 gpurun_out/pytest_rb.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 400 python tools/sweep.py --workload b32m --reps 5 --only 4,5 > gpurun_out/sweep_b32m.log 2>&1; echo "b32m $?"
timeout 400 python tools/sweep.py --workload c5 --reps 5 --only 4,5 > gpurun_out/sweep_c5.log 2>&1; echo "c5 $?"
echo finished
