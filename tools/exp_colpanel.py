#!/usr/bin/env python
"""tools/exp_colpanel.py — column-panel SpMV on the north star's SpMV shape (c3v: uniform-random 16M x 32, fp64):
time per call for P panels (x slice = 134 MB / P) and G lanes per row, against the plain kernels (development aid)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    K = pkg.kernels
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    wl_key = sys.argv[1] if len(sys.argv) > 1 else "c3v"
    wl = bench.WORKLOADS[wl_key]
    td = torch.float64 if wl["dtype"] == "f64" else torch.float32
    esz = 8 if wl["dtype"] == "f64" else 4
    rows, cols, rp, ci, v = bench.gen_matrix_gpu(pkg, wl, td, dev)
    nnz = v.numel()
    x = pkg.synth.fill_uniform(torch.empty(cols, dtype=td, device=dev), stream_id=11)
    y = torch.ones(rows, dtype=td, device=dev)
    ab = bench.algorithmic_bytes("spmv", rows, cols, nnz, 1, esz)
    peak = bench.measured_peaks()[0]

    def timeit(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    plan = K.Plan(rp, nnz).prepare(ci, cols, esz)
    print("col_jump", plan.col_jump(), flush=True)
    base = {}
    for name, o in (("vector_auto", None), ("pipe", pkg.make_opts(variant=pkg.SPMV_PIPE))):
        ms = timeit(lambda: K.spmv(rp, ci, v, x, y, 1.0, 1.0, plan=plan, opts=o, cols=cols))
        base[name] = ms
        print(json.dumps(dict(kernel=name, ms=ms, frac=ab / ms / 1e6 / peak)), flush=True)
    yref = torch.ones(rows, dtype=td, device=dev)
    K.spmv(rp, ci, v, x, yref, 1.0, 1.0, plan=plan, opts=pkg.make_opts(variant=pkg.SPMV_PIPE), cols=cols)
    tiles = [int(a) for a in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["0"])]
    for P, tile in [(int(a), t) for a in (sys.argv[2].split(",") if len(sys.argv) > 2 else "2,3,4,6,8,12,16".split(",")) for t in tiles]:
        if tile:
            os.environ["SBLAS_COLPANEL_TILE"] = str(tile)
        n, W = plan.build_colpanels(ci, v, cols, n_panels=P)
        print("tile", tile, "format(is_sell, stored)", plan.colpanel_format(), "nnz", nnz, flush=True)
        for lanes in (0, 1, 4, 8, 16):
            o = pkg.make_opts(lanes=lanes)
            ms = timeit(lambda: K.spmv(rp, ci, v, x, y, 1.0, 1.0, plan=plan, opts=o, cols=cols))
            y2 = torch.ones(rows, dtype=td, device=dev)
            K.spmv(rp, ci, v, x, y2, 1.0, 1.0, plan=plan, opts=o, cols=cols)
            err = float((y2 - yref).abs().max())
            print(json.dumps(dict(panels=n, panel_cols=W, slice_mb=W * esz / 2**20, lanes=lanes, ms=ms,
                                  frac=ab / ms / 1e6 / peak, max_abs_diff_vs_pipe=err)), flush=True)


if __name__ == "__main__":
    main()
