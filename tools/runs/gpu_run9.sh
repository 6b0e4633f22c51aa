#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -q -x --timeout=60 -p no:cacheprovider -k "spmv" > gpurun_out/pytest_split.log 2>&1
rc=$?; echo "pytest spmv exit $rc"; tail -12 gpurun_out/pytest_split.log
if [ $rc -ne 0 ]; then echo "parity failed: skipping sweeps"; exit 0; fi
timeout 300 python tools/sweep.py --workload c4 --reps 10 --only 3 > gpurun_out/sweep_c4.log 2>&1; echo "c4 $?"
echo finished
