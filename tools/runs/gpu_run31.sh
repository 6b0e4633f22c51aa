#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=200 -p no:cacheprovider -k "gather_heavy or spmv_vs_oracle or golden" > gpurun_out/pytest_gh.log 2>&1
echo "pytest exit $?"; tail -6 gpurun_out/pytest_gh.log
timeout 300 python bench.py --workload c3v --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_c3v.json 2> gpurun_out/bench_c3v.err; echo "c3v $?"; tail -1 gpurun_out/bench_c3v.json | cut -c1-700; tail -2 gpurun_out/bench_c3v.err
timeout 300 python bench.py --workload b32 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_b32.json 2> gpurun_out/bench_b32.err; echo "b32 $?"; tail -1 gpurun_out/bench_b32.json | cut -c1-400
echo finished
