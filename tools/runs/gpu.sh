#!/bin/bash
# tools/runs/gpu.sh STEP [STEP...] — the one script sent to the GPU box (gpurun -- 'bash tools/runs/gpu.sh ...').
# Every step writes under gpurun_out/ (merged back by gpurun); a failing step does not stop the later ones.
#   tests[:EXPR]         pytest -m gpu (optionally -k EXPR)
#   smoke                __graft_entry__.smoke()
#   bench:WL[:ARGS]      python bench.py --workload WL ARGS (one GPU)          -> bench_WL.json
#   benchn:N:WL[:ARGS]   torchrun, N ranks                                     -> bench_WL_nN.json
#   ref[:WL]             python bench.py --impl reference                      -> bench_ref.json
#   launches:WL[:ARGS]   ncu launch list (gpu__time_duration) of bench.py      -> launches_WL.csv
#   ncu:WL:REGEX[:ARGS]  ncu --set full of kernels matching REGEX              -> ncu_WL.ncu-rep (+ summary txt)
#   sweep:WL[:ARGS]      tools/sweep.py                                        -> sweep_WL.log
#   dist:N               tools/check_dist.py under torchrun                    -> check_dist_nN.log
#   multi                tests/test_gpu_multi.py test_dropin.py test_gpu_dist.py on all the devices of the box
#   drv:ARGS             drivers/bin/bench_dropin ARGS                         -> dropin.log (appended)
#   um                   managed-memory drivers, tuning 0 / 1                  -> um.log
#   py:FILE[:ARGS]       python FILE ARGS                                      -> py_<basename>.log
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks.csv 2>/dev/null &
SMI=$!
trap 'kill $SMI 2>/dev/null' EXIT
PORT=29511
run() { echo "=== $* (t=$SECONDS s)"; }
for step in "$@"; do
  IFS=':' read -r kind a b c <<< "$step"
  case "$kind" in
    tests)
      run "$step"
      timeout 2400 python -m pytest tests -m gpu -q -rf --timeout=600 -p no:cacheprovider ${a:+-k "$a"} > gpurun_out/pytest_gpu.log 2>&1
      echo "pytest exit $?"; grep -E "^(FAILED|ERROR)" gpurun_out/pytest_gpu.log | cut -c1-300 | head -40; tail -3 gpurun_out/pytest_gpu.log ;;
    smoke)
      run "$step"
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log ;;
    bench)
      run "$step"
      tag=${a}$(echo "$b" | tr -c 'a-zA-Z0-9\n' '_')
      timeout 900 python bench.py --workload $a $b > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
      echo "bench $a exit $?"; tail -1 gpurun_out/bench_$tag.json | cut -c1-1500; tail -3 gpurun_out/bench_$tag.err ;;
    benchn)
      run "$step"
      tag=${b}_n${a}$(echo "$c" | tr -c 'a-zA-Z0-9\n' '_')
      PORT=$((PORT+1))
      timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $a --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $a --workload $b $c > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
      echo "benchn $a $b exit $?"; grep '^{' gpurun_out/bench_$tag.json | tail -1 | cut -c1-1500; tail -3 gpurun_out/bench_$tag.err ;;
    ref)
      run "$step"
      timeout 600 python bench.py --impl reference ${a:+--workload $a} --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
      echo "ref exit $?"; tail -1 gpurun_out/bench_ref.json | cut -c1-600 ;;
    launches)
      run "$step"
      args="--workload $a --steps 3 --warmup 3 --no-cpu --no-e2e --no-others --no-parity $b"
      timeout 600 python bench.py $args > gpurun_out/launches_$a.plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$a.csv \
        python bench.py $args > gpurun_out/launches_$a.ncu.log 2>&1
      echo "launches $a exit $?" ;;
    ncu)
      run "$step"
      args="--workload $a --steps 2 --warmup 3 --no-cpu --no-e2e --no-others --no-parity $c"
      timeout 600 python bench.py $args > gpurun_out/ncu_$a.plain.log 2>&1 &&
      timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$b" -s 4 -c 2 -f -o gpurun_out/ncu_$a \
        python bench.py $args > gpurun_out/ncu_$a.log 2>&1
      echo "ncu $a exit $?"
      [ -f gpurun_out/ncu_$a.ncu-rep ] && python tools/ncu_summary.py gpurun_out/ncu_$a.ncu-rep gpurun_out/ncu_${a}_summary.txt
      # gpurun brings back at most 64 MiB: keep the summary, drop big reports
      [ -f gpurun_out/ncu_$a.ncu-rep ] && [ $(stat -c %s gpurun_out/ncu_$a.ncu-rep) -gt 12000000 ] && rm -f gpurun_out/ncu_$a.ncu-rep ;;
    sweep)
      run "$step"
      timeout 1200 python tools/sweep.py --workload $a $b > gpurun_out/sweep_$a.log 2>&1; echo "sweep $a exit $?"; tail -30 gpurun_out/sweep_$a.log ;;
    dist)
      run "$step"
      PORT=$((PORT+1))
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $a --master-addr 127.0.0.1 --master-port $PORT \
        tools/check_dist.py > gpurun_out/check_dist_n$a.log 2>&1
      echo "check_dist $a exit $?"; tail -12 gpurun_out/check_dist_n$a.log ;;
    multi)
      run "$step"
      # tests/conftest.py maps logical GPU i to device i % (devices present): the identity on an 8-GPU box
      timeout 1800 python -m pytest tests/test_gpu_multi.py tests/test_dropin.py tests/test_gpu_dist.py \
        -m gpu -q --timeout=900 -p no:cacheprovider -rs > gpurun_out/pytest_multi_devices.log 2>&1
      echo "multi exit $?"; tail -8 gpurun_out/pytest_multi_devices.log ;;
    drv)
      run "$step"
      timeout 900 drivers/bin/bench_dropin $a $b $c >> gpurun_out/dropin.log 2>&1; echo "drv exit $?"; tail -12 gpurun_out/dropin.log ;;
    um)
      run "$step"
      timeout 900 python tools/um_numbers.py > gpurun_out/um.log 2>&1; echo "um exit $?"; tail -30 gpurun_out/um.log ;;
    py)
      run "$step"
      timeout 1500 python $a $b $c > gpurun_out/py_$(basename $a .py).log 2>&1; echo "py $a exit $?"; tail -30 gpurun_out/py_$(basename $a .py).log ;;
    *) echo "unknown step $step" ;;
  esac
done
echo "finished (t=$SECONDS s)"
