#!/bin/bash
# final N-GPU check of the default bench command (what the driver's scaling run launches) + the dist test
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/bench_default_n$N.json 2> gpurun_out/bench_default_n$N.err
echo "bench default n=$N exit $?"; tail -1 gpurun_out/bench_default_n$N.json | cut -c1-300
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_default_n$N.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['config'].get('stitch'), d['config'].get('stitch_requested'), d.get('clocks'))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
echo "ref arm n=$N exit $?"; tail -1 gpurun_out/bench_ref_n$N.json | cut -c1-200
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q -x --timeout=500 -p no:cacheprovider > gpurun_out/pytest_dist_n$N.log 2>&1; echo "pytest dist $?"; tail -2 gpurun_out/pytest_dist_n$N.log
