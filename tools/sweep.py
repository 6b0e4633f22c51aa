#!/usr/bin/env python
"""tools/sweep.py — kernel-variant sweep on one GPU (development aid, not the bench contract).

    python tools/sweep.py --workload c2 [--reps 30] [--scale-down 1]

Times each (variant, tile, lanes, flags) combination with CUDA events, L2 flushed between
launches, and prints median / min ms, GFLOP/s and algorithmic GB/s."""
import argparse
import itertools
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--scale-down", type=int, default=1)
    ap.add_argument("--grid", default="default")
    ap.add_argument("--out", default=None)
    ap.add_argument("--dirty-flush", action="store_true")
    ap.add_argument("--only", default=None, help="only these variants (comma-separated numbers)")
    ap.add_argument("--rb", type=int, default=0, help="row-block height of the SpMM mirror (0 = auto, 4, 8)")
    args = ap.parse_args()
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    K = pkg.kernels
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    wl = bench.scaled(bench.WORKLOADS[args.workload], args.scale_down)
    f64 = wl["dtype"] == "f64"
    td = torch.float64 if f64 else torch.float32
    esz = 8 if f64 else 4
    k = wl.get("k", 1)
    rows, cols, rp, ci, v = bench.gen_matrix_gpu(pkg, wl, td, dev)
    nnz = v.numel()
    plan = K.Plan(rp, nnz)
    print("workload", args.workload, "rows", rows, "nnz", nnz, "k", k, "plan", plan.stats(), flush=True)
    xB = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=td, device=dev), stream_id=11)
    yC = torch.ones(rows * k, dtype=td, device=dev)
    flush = torch.empty(2 * bench.L2_BYTES // 4, dtype=torch.int32, device=dev)
    flush_rd = torch.ones_like(flush)
    ab = bench.algorithmic_bytes(wl["op"], rows, cols, nnz, k, esz)
    if wl["op"] == "spmv":
        grid = [dict(variant=0)]
        grid += [dict(variant=pkg.SPMV_TILE, tile=t, lanes=l) for t in (1024, 1536, 2048, 3072, 4096, 6144, 8192)
                 for l in ((0,) if args.grid == "default" else (0, 1, 2, 4, 8, 16, 32))]
        grid += [dict(variant=pkg.SPMV_SPLIT, tile=t, flags=f) for t in (1536, 2048, 3072, 4096, 6144) for f in (0,)]
        grid += [dict(variant=pkg.SPMV_PIPE, tile=t, stages=st, lanes=l) for t in (1536, 2048, 3072, 4096, 6144, 8192) for st in (2, 3, 4)
                 for l in ((0,) if args.grid == 'default' else (0, 1, 2, 4, 8, 16))]
        grid += [dict(variant=pkg.SPMV_VECTOR, lanes=l, flags=f) for l in (1, 2, 4, 8, 16, 32) for f in (0, 1)]
    else:
        grid = [dict(tile=t, flags=f, lanes=l) for t in (256, 512, 1024, 2048, 4096, 16384) for f in (0, 1)
                for l in ((0,) if args.grid == "default" else (0, 32))]
    if wl['op'] == 'spmm':
        print('rowblocks', plan.build_rowblocks(ci, v, R=args.rb), flush=True)
        grid = [dict(variant=pkg.SPMM_ROWBLOCK_PIPE, flags=f, stages=st, lanes=P) for f in (0, 1) for st in (2,) for P in (1,)] + [dict(variant=pkg.SPMM_ROWBLOCK_TILEB, flags=f) for f in (0, 1)] + [dict(variant=pkg.SPMM_ROWBLOCK, flags=f) for f in (0, 1)] + [dict(variant=pkg.SPMM_ROWMAJOR, tile=t, flags=f) for t in (512, 2048) for f in (0, 1)]
    res = []
    if args.only is not None:
        keep = {int(x) for x in str(args.only).split(",")}
        grid = [g for g in grid if g.get('variant') in keep]
    for g in grid:
        o = pkg.make_opts(**g)

        def run():
            if wl["op"] == "spmv":
                K.spmv(rp, ci, v, xB, yC, 1.0, 1.0, plan=plan, opts=o, cols=cols)
            else:
                K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, xB, k, yC, k, 1.0, 1.0, cols, plan=plan, opts=o)
        try:
            for _ in range(3):
                run()
            torch.cuda.synchronize()
            ts = []
            for _ in range(args.reps):
                flush.zero_()
                if not args.dirty_flush:
                    flush_rd.sum()   # leave L2 full of CLEAN lines (no write-backs billed to the kernel)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                run()
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            med, mn = float(np.median(ts)), float(np.min(ts))
            r = dict(g, ms_med=med, ms_min=mn, gflops=2.0 * nnz * k / med / 1e6, gbs=ab / med / 1e6,
                     frac=ab / med / 1e6 / bench.measured_peaks()[0])
        except Exception as e:  # keep sweeping
            r = dict(g, error=str(e)[:200])
        res.append(r)
        print(json.dumps(r), flush=True)
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
