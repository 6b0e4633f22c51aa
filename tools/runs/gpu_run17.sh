#!/bin/bash
mkdir -p gpurun_out
for pf in 1 2 0; do
  SBLAS_RBP_PREFETCH=$pf timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=90 -p no:cacheprovider -k "rowblock" > gpurun_out/pytest_pf$pf.log 2>&1
  echo "pytest pf=$pf exit $?"; tail -2 gpurun_out/pytest_pf$pf.log
  SBLAS_RBP_PREFETCH=$pf timeout 400 python tools/sweep.py --workload b32m --reps 5 --only 4 > gpurun_out/sweep_b32m_pf$pf.log 2>&1; echo "b32m $?"
  SBLAS_RBP_PREFETCH=$pf timeout 400 python tools/sweep.py --workload c5 --reps 5 --only 4 > gpurun_out/sweep_c5_pf$pf.log 2>&1; echo "c5 $?"
done
echo finished
