#!/bin/bash
# ncu captures of the final kernels (each after the same command exited 0 without ncu);
# summaries are made on the box, only two reports travel back (64 MiB cap on gpurun_out/)
mkdir -p gpurun_out
cap() {  # name workload kernel-regex keep-report
  timeout 300 python bench.py --workload $2 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain_$1.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$3 -s 2 -c 1 -o gpurun_out/prof_$1 -f python bench.py --workload $2 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_$1.log 2>&1
  echo "ncu $1 $?"
  python tools/ncu_summary.py gpurun_out/prof_$1.ncu-rep gpurun_out/r01_ncu_$1_summary.txt
  if [ "$4" != "keep" ]; then rm -f gpurun_out/prof_$1.ncu-rep; fi
}
cap c2_final c2 spmv_pipe keep
cap b32m_final b32m spmm_rbp keep
cap c5_final c5 spmm_rbp drop
cap c3_final c3 spmm_rm drop
cap c4_final c4 spmv_split_kernel drop
# launch list of the default bench command
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/plain_launch.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
echo "launch list $?"
timeout 400 python tools/sweep.py --workload c4 --reps 10 --only 3 > gpurun_out/sweep_c4.log 2>&1; echo "c4 $?"
timeout 400 python tools/sweep.py --workload c5 --reps 5 --only 4 > gpurun_out/sweep_c5.log 2>&1; echo "c5 $?"
du -sh gpurun_out
echo finished
