#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --workload b32m --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/plain_b32m.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rbp -s 1 -c 1 -o gpurun_out/prof_b32m_ring -f python bench.py --workload b32m --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_b32m.log 2>&1
echo "ncu b32m $?"
echo finished
