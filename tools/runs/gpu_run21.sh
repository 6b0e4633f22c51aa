#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
rc=$?; echo "pytest gpu exit $rc"; tail -6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref $?"
timeout 900 python bench.py --workload c5 --no-cpu > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 $?"
timeout 900 python bench.py --workload c4 --no-cpu > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 $?"
timeout 900 python bench.py --workload b32m --no-cpu > gpurun_out/bench_b32m.json 2> gpurun_out/bench_b32m.err; echo "bench b32m $?"
timeout 900 python bench.py --workload b32m --no-cpu --flags 1 > gpurun_out/bench_b32m_fma.json 2> gpurun_out/bench_b32m_fma.err; echo "bench b32m fma $?"
timeout 900 python bench.py --workload c3v --no-cpu > gpurun_out/bench_c3v.json 2> gpurun_out/bench_c3v.err; echo "bench c3v $?"
echo finished
