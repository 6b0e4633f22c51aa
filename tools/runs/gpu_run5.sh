#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=180 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all $?"; tail -25 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench exit $?"; cat gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
timeout 300 python tools/sweep.py --workload c2 --reps 20 --only 4 > gpurun_out/sweep_c2.log 2>&1; echo "c2 $?"
echo finished
