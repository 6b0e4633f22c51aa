#!/bin/bash
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload c4 --scaling strong --steps 100 --warmup 5 --no-cpu --stitch mc > gpurun_out/bench_c4s_mc_n$N.json 2> gpurun_out/bench_c4s_mc_n$N.err
echo "exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c4s_mc_n$N.json').read().strip().splitlines()[-1])
print('   value %.1f ms %.4f kernel_ms %.4f stitch %s mc %s' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['config'].get('stitch'), d['config'].get('multicast')))
PY
tail -2 gpurun_out/bench_c4s_mc_n$N.err | cut -c1-300
