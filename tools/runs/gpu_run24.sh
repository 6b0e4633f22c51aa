#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --workload b32m --variant 5 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain_b32m_t.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rbt -s 2 -c 1 -o gpurun_out/prof_b32m_tileb -f python bench.py --workload b32m --variant 5 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_b32m_t.log 2>&1
echo "ncu $?"
python tools/ncu_summary.py gpurun_out/prof_b32m_tileb.ncu-rep gpurun_out/r01_ncu_b32m_tileb_summary.txt
echo finished
