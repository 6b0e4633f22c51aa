#!/bin/bash
# usage: gpu_multiN.sh N   — config 4 (R-MAT fp32 SpMV, nnz-balanced panels) strong and weak, config 5 at N GPUs
N=$1
mkdir -p gpurun_out
run() {  # tag, bench args...
  tag=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_n$N.json 2> gpurun_out/bench_${tag}_n$N.err
  echo "bench $tag n=$N exit $?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${tag}_n$N.json').read().strip().splitlines()[-1])
    print('   value %.1f ms %.4f kernel_ms %.4f floor %.4f stitch %s scaling %s' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['config'].get('nvlink_floor_ms_per_step', 0), d['config'].get('stitch'), d['scaling']))
except Exception as e:
    print('   ERR', e)
PY
  tail -2 gpurun_out/bench_${tag}_n$N.err | cut -c1-300
}
run c4s --workload c4 --scaling strong --steps 100 --warmup 5 --no-cpu
run c4w --workload c4 --scaling weak --steps 50 --warmup 5 --no-cpu
if [ "$N" = "8" ]; then run c5 --workload c5 --no-cpu; fi
echo finished
