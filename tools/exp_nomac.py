#!/usr/bin/env python
"""experiment: time the row-block SpMM kernels with the block loop skipped (flag bit 30 in a
scratch build): what is left is the data movement of the pipeline (A stage, C in/out, B staging)."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("s-blas_b200")
K = pkg.kernels
dev = torch.device("cuda:0")
wl = bench.WORKLOADS["b32m"]
rows, cols, rp, ci, v = bench.gen_matrix_gpu(pkg, wl, torch.float64, dev)
k = 64
plan = K.Plan(rp, v.numel()); print(plan.build_rowblocks(ci, v))
B = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=torch.float64, device=dev), stream_id=11)
Cm = torch.ones(rows * k, dtype=torch.float64, device=dev)
for var in (4, 5):
    for fl in (0, 1 << 30, 1, 1 | (1 << 30)):
        o = pkg.make_opts(variant=var, flags=fl)
        for _ in range(3):
            K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, B, k, Cm, k, 1.0, 1.0, cols, plan=plan, opts=o)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, B, k, Cm, k, 1.0, 1.0, cols, plan=plan, opts=o)
        b.record(); torch.cuda.synchronize()
        print("variant", var, "flags", hex(fl), "ms/step %.3f" % (a.elapsed_time(b) / 10), flush=True)
