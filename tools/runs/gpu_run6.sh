#!/bin/bash
# Multi-GPU bring-up (run with gpurun --gpus N): single-process mgpu ops on real peers, then the
# torchrun bench at N ranks for each multi-GPU config.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$N.txt; nvidia-smi topo -m >> gpurun_out/gpus_$N.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_dropin.py -m gpu -q --timeout=180 -p no:cacheprovider > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest multi $?"; tail -5 gpurun_out/pytest_multi_$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for wl in c2 c4; do
  for st in peer nccl; do
    timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 --workload $wl --stitch $st --no-cpu > gpurun_out/bench_${wl}_${st}_n$N.json 2> gpurun_out/bench_${wl}_${st}_n$N.err; echo "$wl $st $?"; cat gpurun_out/bench_${wl}_${st}_n$N.json | cut -c1-600; tail -3 gpurun_out/bench_${wl}_${st}_n$N.err
  done
done
for st in peer nccl; do
  timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload c3 --stitch $st --no-cpu --no-e2e > gpurun_out/bench_c3_${st}_n$N.json 2> gpurun_out/bench_c3_${st}_n$N.err; echo "c3 $st $?"; cat gpurun_out/bench_c3_${st}_n$N.json | cut -c1-600; tail -3 gpurun_out/bench_c3_${st}_n$N.err
done
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload c5 --no-cpu --no-e2e > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 $?"; cat gpurun_out/bench_c5_n$N.json | cut -c1-600; tail -3 gpurun_out/bench_c5_n$N.err
echo finished
