#!/usr/bin/env python
"""SURVEY §8(f) f4 — the "reference-GPU" comparison lane (measurement tooling, not product code).

For one bench workload on ONE B200 it times, on the SAME synthetic operands:
  ours      our C-ABI kernel (resident operands, CUDA events, back-to-back steps — bench.py's protocol)
  cusparse  the cuSPARSE call the reference issues (spmv.h:70-104 / spmm.h:199-250: CSR int32,
            ALG_DEFAULT, column-major B / C), resident operands, CUDA events around the call alone
  ref_op    the UNMODIFIED reference op (sblas_spmv_csr_v1 / sblas_spmm_csr_v1 / _v2, compiled from
            /root/reference into oracle/_ref/libsblas_refgpu.so) as its caller sees it: NCCL communicator
            and cuSPARSE handle creation, zero-filled C_copy upload, all-reduce and axpby are inside the
            call, as in the reference's own timed region (spmv_test.cu:20-27)
and checks the three outputs against each other (per-element |d| <= rel * (|A||x| + |y|) proxy: max
abs difference against the output scale).  One JSON line per workload on stdout.

    python tools/ref_gpu_lane.py --workload c2 [--reps 5] [--scale-down 1]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B          # workloads and generators only  # noqa: E402

REFGPU_SO = os.path.join(ROOT, "oracle", "_ref", "libsblas_refgpu.so")


class quiet_stdout:
    """the reference prints per-GPU lines from inside its ops: park fd 1 on /dev/null meanwhile"""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        nul = os.open(os.devnull, os.O_WRONLY)
        os.dup2(nul, 1)
        os.close(nul)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.saved)


def _p(t):
    return C.c_void_p(t.data_ptr())


def _np(a):
    return a.ctypes.data_as(C.c_void_p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2", choices=list(B.WORKLOADS))
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--ref-reps", type=int, default=2)
    ap.add_argument("--scale-down", type=int, default=1)
    ap.add_argument("--no-ref-op", action="store_true")
    args = ap.parse_args()

    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    K = pkg.kernels
    if not os.path.exists(REFGPU_SO):
        raise SystemExit("oracle/_ref/libsblas_refgpu.so missing: run `make -C oracle refgpu` where /root/reference exists")
    R = C.CDLL(REFGPU_SO)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    wl = B.scaled(B.WORKLOADS[args.workload], args.scale_down)
    op, f64 = wl["op"], wl["dtype"] == "f64"
    tdt = torch.float64 if f64 else torch.float32
    esz = 8 if f64 else 4
    k = wl.get("k", 1)
    method = wl.get("method", 2)
    rows, cols, rp, ci, v = B.gen_matrix_gpu(pkg, wl, tdt, dev)
    nnz = v.numel()
    flops = 2.0 * nnz * k
    abytes = B.algorithmic_bytes(op, rows, cols, nnz, k, esz)
    out = {"workload": args.workload + ": " + wl["name"], "rows": rows, "cols": cols, "nnz": nnz, "k": k,
           "dtype": wl["dtype"], "n_gpus": 1, "algorithmic_bytes": abytes}

    def ev_time(fn, reps, warm=2):
        for _ in range(warm):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    # ---------------- ours (row-major device operands as sync2gpu lays them out)
    plan = K.Plan(rp, nnz)
    if op == "spmm":
        plan.build_rowblocks(ci, v)
    else:   # what CsrSparseMatrix::sync2gpu does for SpMV: tile table, col_jump, column-panel mirror when it pays
        plan.prepare(ci, cols, esz)
        out["colpanels"] = plan.build_colpanels(ci, v, cols)
    xB = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=tdt, device=dev), stream_id=11)
    yC = torch.ones(rows * k, dtype=tdt, device=dev)
    if op == "spmv":
        ours = lambda: K.spmv(rp, ci, v, xB, yC, 1.0, 1.0, plan=plan, cols=cols)      # noqa: E731
    else:
        ours = lambda: K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, xB, k, yC, k, 1.0, 1.0, cols, plan=plan)  # noqa: E731
    ms_ours = ev_time(ours, args.reps * (20 if op == "spmv" else 1))
    yC.fill_(1.0)
    ours()
    torch.cuda.synchronize()
    out["ours"] = {"ms": ms_ours, "gflops": flops / ms_ours / 1e6, "gbs": abytes / ms_ours / 1e6}
    del plan

    # ---------------- cuSPARSE on resident operands; col-major B / C as the reference keeps them
    if op == "spmm":
        Bcm = torch.empty(cols * k, dtype=tdt, device=dev)
        K.transpose(xB, k, cols, k, Bcm, cols)          # row-major [cols x k] -> col-major
        ours_cm = torch.empty(rows * k, dtype=tdt, device=dev)
        K.transpose(yC, k, rows, k, ours_cm, rows)
        torch.cuda.synchronize()
        del yC, xB
        torch.cuda.empty_cache()
        Ccm = torch.ones(rows * k, dtype=tdt, device=dev)
    else:
        Bcm, ours_cm = xB, yC
        Ccm = torch.ones(rows, dtype=tdt, device=dev)
    nrep = args.reps * (20 if op == "spmv" else 1)
    ms = (C.c_float * nrep)()
    # beta = 0 for the timing loop would change the traffic; keep alpha = beta = 1 and accept that C
    # accumulates over the repetitions, then redo ONE call on a fresh C for the comparison
    if op == "spmv":
        call = lambda warm, reps: R.libgpu_cusparse_spmv(  # noqa: E731
            C.c_int(int(f64)), C.c_long(rows), C.c_long(cols), C.c_long(nnz), _p(rp), _p(ci), _p(v), _p(Bcm),
            _p(Ccm), C.c_double(1.0), C.c_double(1.0), C.c_int(warm), C.c_int(reps), ms)
    else:
        call = lambda warm, reps: R.libgpu_cusparse_spmm(  # noqa: E731
            C.c_int(int(f64)), C.c_long(rows), C.c_long(cols), C.c_long(nnz), _p(rp), _p(ci), _p(v), C.c_long(k),
            _p(Bcm), _p(Ccm), C.c_double(1.0), C.c_double(1.0), C.c_int(warm), C.c_int(reps), ms)
    rc = call(2, nrep)
    if rc != 0:
        out["cusparse"] = {"error": rc}
    else:
        t = float(np.median(np.array(ms[:nrep])))
        Ccm.fill_(1.0)
        call(0, 1)
        torch.cuda.synchronize()
        scale = float(ours_cm.abs().max())
        dmax = float((Ccm - ours_cm).abs().max())
        out["cusparse"] = {"ms": t, "gflops": flops / t / 1e6, "gbs": abytes / t / 1e6,
                           "max_abs_diff_vs_ours": dmax, "output_scale": scale,
                           "version": "libcusparse of CUDA 12.9 (image)", "alg": "ALG_DEFAULT, CSR, col-major B/C"}
        out["speedup_vs_cusparse_kernel"] = t / ms_ours

    # ---------------- the reference op itself, host operands (its sync2gpu uploads them)
    if not args.no_ref_op:
        rp_h, ci_h, v_h = rp.cpu().numpy(), ci.cpu().numpy(), v.cpu().numpy()
        B_h = Bcm.cpu().numpy()
        ours_h = ours_cm.cpu().numpy()
        del Bcm, Ccm, ours_cm, rp, ci, v
        torch.cuda.empty_cache()
        C0 = np.ones(rows * k, v_h.dtype)
        Cout = np.empty(rows * k, v_h.dtype)
        msr = (C.c_double * args.ref_reps)()
        with quiet_stdout():
            if op == "spmv":
                fn = R.refgpu_spmv_f64 if f64 else R.refgpu_spmv_f32
                ct = C.c_double if f64 else C.c_float
                rc = fn(C.c_int(rows), C.c_int(cols), C.c_int(nnz), _np(rp_h), _np(ci_h), _np(v_h), _np(B_h), _np(C0),
                        _np(Cout), ct(1.0), ct(1.0), C.c_uint(1), C.c_int(args.ref_reps), msr)
            elif f64:
                rc = R.refgpu_spmm_f64(C.c_int(method), C.c_int(rows), C.c_int(cols), C.c_int(nnz), _np(rp_h),
                                       _np(ci_h), _np(v_h), C.c_int(k), _np(B_h), _np(C0), _np(Cout), C.c_double(1.0),
                                       C.c_double(1.0), C.c_uint(1), C.c_int(args.ref_reps), msr)
            else:
                rc = -9
        if rc != 0:
            out["ref_op"] = {"error": rc}
        else:
            ts = [float(x) for x in msr]
            t = min(ts)
            out["ref_op"] = {"ms_per_call": ts, "ms": t, "gflops": flops / t / 1e6,
                             "entry": {"spmv": "sblas_spmv_csr_v1 spmv.h:36", "spmm": "sblas_spmm_csr_v%d" % method}[op],
                             "includes": "ncclCommInitRank, cusparseCreate, zero C_copy alloc + H2D, cuSPARSE, "
                                         "ncclAllReduce, axpby (all inside the reference's call)",
                             "max_abs_diff_vs_ours": float(np.abs(Cout - ours_h).max())}
            out["speedup_vs_ref_op"] = t / ms_ours
    print(json.dumps(out))


if __name__ == "__main__":
    main()
