#!/bin/bash
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=60 -p no:cacheprovider -k "spmv" > gpurun_out/pytest_pipe.log 2>&1
rc=$?; echo "pytest spmv exit $rc" >> gpurun_out/pytest_pipe.log; tail -15 gpurun_out/pytest_pipe.log
if [ $rc -ne 0 ]; then echo "parity failed: skipping sweeps"; exit 0; fi
timeout 300 python tools/sweep.py --workload c2 --reps 20 --only 4 > gpurun_out/sweep_c2.log 2>&1; echo "c2 $?"
timeout 300 python tools/sweep.py --workload b32 --reps 10 --only 4 --grid full > gpurun_out/sweep_b32.log 2>&1; echo "b32 $?"
timeout 300 python tools/sweep.py --workload c3v --reps 5 --only 4 > gpurun_out/sweep_c3v.log 2>&1; echo "c3v $?"
echo finished
