#!/bin/bash
# First GPU pass: parity tests, smoke, bench, ncu launch list.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench exit $?"; cat gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
timeout 600 python tools/sweep.py --workload c2 --reps 20 > gpurun_out/sweep_c2.log 2>&1; tail -40 gpurun_out/sweep_c2.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_list.log 2>&1
echo finished
