#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
rc=$?; echo "pytest gpu exit $rc"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke $?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 $?"
timeout 900 python bench.py --workload b32m --no-cpu > gpurun_out/bench_b32m.json 2> gpurun_out/bench_b32m.err; echo "bench b32m $?"
timeout 900 python bench.py --workload c5 --no-cpu > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 $?"
timeout 900 python bench.py --workload c5 --no-cpu --no-e2e --k 32 --variant 1 > gpurun_out/bench_c5_k32_v1.json 2> gpurun_out/bench_c5_k32_v1.err; echo "bench c5 k32 v1 $?"
timeout 900 python bench.py --workload c5 --no-cpu --no-e2e --k 32 --variant 4 > gpurun_out/bench_c5_k32_v4.json 2> gpurun_out/bench_c5_k32_v4.err; echo "bench c5 k32 v4 $?"
timeout 900 python bench.py --workload c5 --no-cpu --no-e2e --k 64 > gpurun_out/bench_c5_k64.json 2> gpurun_out/bench_c5_k64.err; echo "bench c5 k64 $?"
echo finished
