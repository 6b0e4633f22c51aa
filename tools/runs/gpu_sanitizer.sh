#!/bin/bash
# compute-sanitizer memcheck over the C++ drivers (ash85, one GPU): SpMV, SpMM method 1 and 2
mkdir -p gpurun_out
A=tests/golden/ash85.mtx
{
  timeout 20 compute-sanitizer --tool memcheck --error-exitcode 9 drivers/bin/spmv_test $A 1.0 1.0 1; echo "spmv rc $?"
  timeout 25 compute-sanitizer --tool memcheck --error-exitcode 9 drivers/bin/spmm_test 2 $A 64 1.0 1.0 1; echo "spmm rc $?"
} > gpurun_out/sanitizer_drivers.log 2>&1
grep -E "rc |ERROR SUMMARY|Validation" gpurun_out/sanitizer_drivers.log
