#!/bin/bash
# PIPE kernel bring-up: parity first (short timeout: a wedged mbarrier must not hold the box), then sweeps.
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=60 -p no:cacheprovider -k "spmv" > gpurun_out/pytest_pipe.log 2>&1
rc=$?; echo "pytest spmv exit $rc" >> gpurun_out/pytest_pipe.log; tail -15 gpurun_out/pytest_pipe.log
if [ $rc -ne 0 ]; then echo "parity failed: skipping sweeps"; exit 0; fi
timeout 300 python tools/sweep.py --workload c2 --reps 20 > gpurun_out/sweep_c2.log 2>&1; echo "c2 $?"
timeout 300 python tools/sweep.py --workload b32 --reps 10 > gpurun_out/sweep_b32.log 2>&1; echo "b32 $?"
timeout 300 python tools/sweep.py --workload c3v --reps 5 > gpurun_out/sweep_c3v.log 2>&1; echo "c3v $?"
timeout 300 python tools/sweep.py --workload c4 --reps 10 > gpurun_out/sweep_c4.log 2>&1; echo "c4 $?"
timeout 600 python -m pytest tests -m gpu -q --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all $?"; tail -3 gpurun_out/pytest_gpu.log
echo finished
