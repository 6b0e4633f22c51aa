#!/bin/bash
# Second GPU pass: parity tests, variant sweeps on every workload, ncu full capture of the c2 SpMV kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep.py --workload b32 --reps 10 > gpurun_out/sweep_b32.log 2>&1; echo "b32 $?"
timeout 600 python tools/sweep.py --workload c3v --reps 10 > gpurun_out/sweep_c3v.log 2>&1; echo "c3v $?"
timeout 600 python tools/sweep.py --workload c4 --reps 10 > gpurun_out/sweep_c4.log 2>&1; echo "c4 $?"
timeout 900 python tools/sweep.py --workload c3 --reps 3 > gpurun_out/sweep_c3.log 2>&1; echo "c3 $?"
timeout 900 python tools/sweep.py --workload c5 --reps 3 > gpurun_out/sweep_c5.log 2>&1; echo "c5 $?"
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmv_tile -s 3 -c 2 -o gpurun_out/prof_c2_spmv -f python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu $?"
echo finished
