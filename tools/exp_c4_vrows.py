#!/usr/bin/env python
"""tools/exp_c4_vrows.py — feasibility probe (development aid): the sliced-ELL per-lane kernel on the R-MAT input
(config 4) with long rows cut into VIRTUAL rows of at most L entries (a second tiny kernel would add the pieces of a
row).  Prints the sliced-ELL padding and the time of the SpMV over the virtual matrix against AUTO (SPLIT)."""
import json
import os
import sys

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
    wl = bench.WORKLOADS["c4"]
    rows, cols, rp, ci, v = bench.gen_matrix_gpu(pkg, wl, torch.float32, dev)
    nnz = v.numel()
    x = pkg.synth.fill_uniform(torch.empty(cols, dtype=torch.float32, device=dev), stream_id=11)

    def timeit(fn, reps=50):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    y = torch.ones(rows, dtype=torch.float32, device=dev)
    plan = K.Plan(rp, nnz).prepare(ci, cols, 4)
    print(json.dumps(dict(kernel="auto (SPLIT)", ms=timeit(lambda: K.spmv(rp, ci, v, x, y, 1.0, 0.0, plan=plan, cols=cols)))), flush=True)
    lens = (rp[1:] - rp[:-1]).to(torch.int64)
    for L in (16, 32, 64, 128):
        for sort_rows in (False, True):
            nch = (lens + L - 1) // L                      # pieces per row (0 for empty rows)
            nch1 = torch.clamp(nch, min=1)                 # keep empty rows as one empty virtual row
            first = torch.cumsum(nch1, 0) - nch1           # first virtual row of each row
            nv = int(nch1.sum())
            owner = torch.repeat_interleave(torch.arange(rows, device=dev), nch1)
            piece = torch.arange(nv, device=dev) - first[owner]
            vlen = torch.clamp(lens[owner] - piece * L, min=0, max=L)
            if sort_rows:   # global sort of the virtual rows by length (what a sigma = infinity window would do)
                order = torch.argsort(vlen, descending=True, stable=True)
                start = rp[:-1].to(torch.int64)[owner] + piece * L
                vlen_s = vlen[order]
                vrp = torch.zeros(nv + 1, dtype=torch.int32, device=dev)
                vrp[1:] = torch.cumsum(vlen_s, 0).to(torch.int32)
                # gather entries into the new order
                src = torch.repeat_interleave(start[order], vlen_s) + (torch.arange(nnz, device=dev) - torch.repeat_interleave(vrp[:-1].to(torch.int64), vlen_s))
                ci2, v2 = ci[src].contiguous(), v[src].contiguous()
            else:
                vrp = torch.zeros(nv + 1, dtype=torch.int32, device=dev)
                vrp[1:] = torch.cumsum(vlen, 0).to(torch.int32)
                ci2, v2 = ci, v
            yv = torch.ones(nv, dtype=torch.float32, device=dev)
            p2 = K.Plan(vrp, nnz)
            n, W = p2.build_colpanels(ci2, v2, cols, n_panels=1)
            is_sell, stored = p2.colpanel_format()
            ms = timeit(lambda: K.spmv(vrp, ci2, v2, x, yv, 1.0, 0.0, plan=p2, cols=cols))
            print(json.dumps(dict(L=L, sorted=sort_rows, virtual_rows=nv, is_sell=is_sell, padding=stored / nnz - 1.0, ms=ms)), flush=True)
            del p2, yv, vrp


if __name__ == "__main__":
    main()
