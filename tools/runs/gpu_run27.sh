#!/bin/bash
# f4: reference-GPU / cuSPARSE comparison lane on one B200
mkdir -p gpurun_out
for w in c2 b32 b32m c5 c4; do
  timeout 600 python tools/ref_gpu_lane.py --workload $w > gpurun_out/lane_$w.json 2> gpurun_out/lane_$w.err
  echo "lane $w exit $?"; tail -c 1500 gpurun_out/lane_$w.json; tail -3 gpurun_out/lane_$w.err
done
echo finished
