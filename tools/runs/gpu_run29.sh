#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain_c4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmv_split_kernel -s 3 -c 1 -o gpurun_out/prof_c4_ovl -f python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_c4.log 2>&1
echo "ncu $?"; tail -2 gpurun_out/plain_c4.log | cut -c1-400
python tools/ncu_summary.py gpurun_out/prof_c4_ovl.ncu-rep gpurun_out/r01_ncu_c4_ovl_summary.txt
cat gpurun_out/r01_ncu_c4_ovl_summary.txt
echo finished
