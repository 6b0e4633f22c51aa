#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 python bench.py --steps 200 --warmup 20 > gpurun_out/bench_c2_n1.json 2> gpurun_out/bench_c2_n1.err; echo "c2 n1 $?"; cut -c1-300 gpurun_out/bench_c2_n1.json; tail -3 gpurun_out/bench_c2_n1.err
for wl in c2 c4; do
  for st in peer nccl; do
    timeout 300 $TR bench.py --gpus $N --steps 100 --warmup 10 --workload $wl --stitch $st --no-cpu > gpurun_out/bench_${wl}_${st}_n$N.json 2> gpurun_out/bench_${wl}_${st}_n$N.err; echo "$wl $st $?"; cut -c1-300 gpurun_out/bench_${wl}_${st}_n$N.json; tail -3 gpurun_out/bench_${wl}_${st}_n$N.err | grep -v OMP_NUM | grep -v "\*\*\*"
  done
done
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload c3 --stitch peer --no-cpu --no-e2e > gpurun_out/bench_c3_peer_n$N.json 2> gpurun_out/bench_c3_peer_n$N.err; echo "c3 peer $?"; cut -c1-300 gpurun_out/bench_c3_peer_n$N.json
echo finished
