#!/bin/bash
# tools/gpu_multi2.sh N — stitch check + stitch comparison on N GPUs
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/check_dist.py > gpurun_out/check_dist_n$N.log 2>&1
echo "check_dist exit $?"; grep -v "^$" gpurun_out/check_dist_n$N.log | tail -16
run() {  # tag, bench args...
  tag=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_n$N.json 2> gpurun_out/bench_${tag}_n$N.err
  echo "bench $tag n=$N exit $?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${tag}_n$N.json').read().strip().splitlines()[-1])
    print('   value %.1f ms %.4f kernel_ms %.4f floor %.4f stitch %s mc %s' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['config'].get('nvlink_floor_ms_per_step', 0), d['config'].get('stitch'), d['config'].get('multicast')))
except Exception as e:
    print('   ERR', e)
PY
}
run c2mc --stitch mc --no-cpu
run c2 --no-cpu
run b32mmc --workload b32m --stitch mc --no-cpu
run b32m --workload b32m --no-cpu
run c3mc --workload c3 --stitch mc --no-cpu
echo finished
