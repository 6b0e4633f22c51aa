#!/bin/bash
# tools/gpu_multi.sh N — the multi-GPU legs on one box with N GPUs
N=${1:-2}
mkdir -p gpurun_out
run() {  # tag, bench args...
  tag=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_n$N.json 2> gpurun_out/bench_${tag}_n$N.err
  echo "bench $tag n=$N exit $?"; tail -c 600 gpurun_out/bench_${tag}_n$N.json | head -c 300; echo
}
if [ "$2" == "tests" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_dropin.py -m gpu -q -x --timeout=300 -p no:cacheprovider > gpurun_out/pytest_multi_n$N.log 2>&1
  echo "pytest multi exit $?"; tail -3 gpurun_out/pytest_multi_n$N.log
fi
run c2 
run c2nccl --stitch nccl
run b32m --workload b32m --no-cpu
run c3 --workload c3 --no-cpu
run c5 --workload c5 --no-cpu
run b32 --workload b32 --no-cpu
echo finished
