#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/calib.py --reps 10 > gpurun_out/calib.log 2>&1; echo "calib $?"; cat gpurun_out/calib.log
timeout 300 python tools/sweep.py --workload c2 --reps 20 > gpurun_out/sweep_c2.log 2>&1; echo "c2 $?"
timeout 300 python tools/sweep.py --workload c2 --reps 20 --dirty-flush --only 4 > gpurun_out/sweep_c2_dirty.log 2>&1; echo "c2d $?"
timeout 300 python tools/sweep.py --workload b32 --reps 10 > gpurun_out/sweep_b32.log 2>&1; echo "b32 $?"
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --variant 4 --tile 1536 > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -s 3 -c 1 -o gpurun_out/prof_c2_pipe -f python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --variant 4 --tile 1536 > gpurun_out/ncu_full.log 2>&1
echo "ncu $?"
echo finished
