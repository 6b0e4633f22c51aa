#!/usr/bin/env python
"""Experiment (development aid): does relabelling the columns of a power-law matrix by descending
frequency (hot x entries contiguous -> L1 sectors hold 8 hot values instead of 1) speed up the SPLIT
SpMV?  The permutation is built with torch here; the kernel is the product's, run on the relabelled
col_idx and the permuted x.   python tools/exp_colperm.py [--workload c4]"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    K = pkg.kernels
    dev = torch.device("cuda", 0)
    wl = bench.WORKLOADS[args.workload]
    f64 = wl["dtype"] == "f64"
    td = torch.float64 if f64 else torch.float32
    rows, cols, rp, ci, v = bench.gen_matrix_gpu(pkg, wl, td, dev)
    nnz = v.numel()
    plan = K.Plan(rp, nnz)
    x = pkg.synth.fill_uniform(torch.empty(cols, dtype=td, device=dev), stream_id=11)
    cnt = torch.bincount(ci.long(), minlength=cols)
    srt = torch.sort(cnt, descending=True, stable=True)
    order = srt.indices                       # new -> old
    cs = torch.cumsum(srt.values, 0).double() / nnz
    for h in (1024, 4096, 16384, 32768, 65536, 262144, 1 << 20):
        print("top %8d columns hold %.3f of nnz" % (h, float(cs[h - 1])))
    perm = torch.empty_like(order)
    perm[order] = torch.arange(cols, device=dev)
    ci2 = perm[ci.long()].int().contiguous()
    x2 = x[order].contiguous()
    flush = torch.empty(2 * bench.L2_BYTES // 4, dtype=torch.int32, device=dev)
    flush_rd = torch.ones_like(flush)

    def timeit(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(args.reps):
            flush.zero_(); flush_rd.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))
    # hot-only permutation: the H hottest columns first (by frequency), the rest in original order
    def hot_perm(H):
        hot = order[:H]
        is_hot = torch.zeros(cols, dtype=torch.bool, device=dev)
        is_hot[hot] = True
        cold = torch.nonzero(~is_hot).squeeze(1)
        o2 = torch.cat([hot, cold])
        p2 = torch.empty_like(o2)
        p2[o2] = torch.arange(cols, device=dev)
        return o2, p2
    for tile in (1024, 1536, 2048):
        for fl in (0, pkg.FLAG_SPLIT_MERGE):
            o = pkg.make_opts(variant=pkg.SPMV_SPLIT, tile=tile, flags=fl)
            y = torch.ones(rows, dtype=td, device=dev)
            t0 = timeit(lambda: K.spmv(rp, ci, v, x, y, 1.0, 1.0, plan=plan, opts=o, cols=cols))
            t1 = timeit(lambda: K.spmv(rp, ci2, v, x2, y, 1.0, 1.0, plan=plan, opts=o, cols=cols))
            print(json.dumps(dict(tile=tile, flags=fl, ms_orig=t0, ms_fullsort=t1)), flush=True)
    o = pkg.make_opts(variant=pkg.SPMV_SPLIT, tile=1536, flags=pkg.FLAG_SPLIT_MERGE)
    for H in (4096, 32768, 262144, 1 << 21):
        o2, p2 = hot_perm(H)
        ci3 = p2[ci.long()].int().contiguous()
        x3 = x[o2].contiguous()
        y = torch.ones(rows, dtype=td, device=dev)
        t = timeit(lambda: K.spmv(rp, ci3, v, x3, y, 1.0, 1.0, plan=plan, opts=o, cols=cols))
        print(json.dumps(dict(hot=H, ms=t)), flush=True)
    # correctness of the relabelling: same y bit for bit
    ya = torch.ones(rows, dtype=td, device=dev); yb = torch.ones(rows, dtype=td, device=dev)
    K.spmv(rp, ci, v, x, ya, 1.0, 1.0, plan=plan, opts=o, cols=cols)
    K.spmv(rp, ci2, v, x2, yb, 1.0, 1.0, plan=plan, opts=o, cols=cols)
    torch.cuda.synchronize()
    print("bit-identical:", bool(torch.equal(ya, yb)))
    tp = timeit(lambda: torch.index_select(x, 0, order))
    print("x permute (torch index_select) ms:", tp)


if __name__ == "__main__":
    main()
