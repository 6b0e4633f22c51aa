#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=120 -p no:cacheprovider -k "spmm" > gpurun_out/pytest_rb.log 2>&1
rc=$?; echo "pytest spmm exit $rc"; tail -15 gpurun_out/pytest_rb.log
if [ $rc -ne 0 ]; then echo "parity failed: skipping sweeps"; exit 0; fi
timeout 600 python tools/sweep.py --workload c5 --reps 3 > gpurun_out/sweep_c5.log 2>&1; echo "c5 $?"; grep -v "^{" gpurun_out/sweep_c5.log | tail -3
timeout 600 python tools/sweep.py --workload b32m --reps 3 > gpurun_out/sweep_b32m.log 2>&1; echo "b32m $?"; grep -v "^{" gpurun_out/sweep_b32m.log | tail -3
timeout 900 python -m pytest tests -m gpu -q --timeout=180 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all $?"; tail -4 gpurun_out/pytest_gpu.log
echo finished
