#!/bin/bash
# balanced merge-path phase 2 of the SPLIT SpMV: parity, then the config-4 sweep against the old phase 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=120 -p no:cacheprovider -k "spmv" > gpurun_out/pytest_spmv.log 2>&1
echo "pytest spmv exit $?"; tail -5 gpurun_out/pytest_spmv.log
timeout 600 python tools/sweep.py --workload c4 --reps 10 --only 3 > gpurun_out/sweep_c4_merge.log 2>&1; echo "c4 $?"
cat gpurun_out/sweep_c4_merge.log | cut -c1-200
echo finished
