#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=60 -p no:cacheprovider -k "spmv" > gpurun_out/pytest_split.log 2>&1
rc=$?; echo "pytest spmv exit $rc"; tail -5 gpurun_out/pytest_split.log
timeout 300 python tools/sweep.py --workload c4 --reps 10 --only 3 > gpurun_out/sweep_c4.log 2>&1; echo "c4 $?"
timeout 300 python bench.py --workload b32m --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/plain_b32m.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rbp -s 1 -c 1 -o gpurun_out/prof_b32m_rbp -f python bench.py --workload b32m --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_b32m.log 2>&1
echo "ncu b32m $?"
timeout 300 python bench.py --workload c5 --variant 3 --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/plain_c5.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rb_kernel -s 1 -c 1 -o gpurun_out/prof_c5_rb -f python bench.py --workload c5 --variant 3 --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_c5.log 2>&1
echo "ncu c5 $?"
echo finished
