#!/usr/bin/env python
"""oracle/gen_golden.py — TEST INFRASTRUCTURE.

Generates tests/golden/* from the UNMODIFIED reference (oracle/_ref/libsblas_ref.so,
built from /root/reference by oracle/Makefile).  Run in the build container:

    make -C oracle && python oracle/gen_golden.py

The vectors are committed; /root/reference does not exist on the GPU box, so tests
read only the committed files.  Cases mirror the reference's own tests:
  * ash85.mtx, x = y = 1, alpha = beta = 1                (spmv_test.cu:12-14, README.md:76)
  * ash85.mtx, alpha = 3, beta = 4, k = 256, 4 GPUs       (unit_test.cu:183-185)
  * ash85.mtx, SpMM k = 64, B = DenseMatrix(w,64,col_major) seed 211, C = 1
  * partition tables for n_gpu in {1,2,3,4,8}            (matrix.h:360-375)
plus seeded random CSR matrices (empty rows, unsorted columns, one long row) in fp64 and
fp32 so that kernel variants other than the short-row one are pinned as well.
"""
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import pyoracle  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF_DIR = "/root/reference"


def rand_csr(rng, h, w, avg, empty_frac=0.2, long_row=None, dtype=np.float64, sort=False):
    lens = rng.poisson(avg, h).astype(np.int64)
    lens[rng.random(h) < empty_frac] = 0
    if long_row is not None:
        lens[long_row[0]] = long_row[1]
    lens = np.minimum(lens, w)
    rp = np.zeros(h + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    ci = np.zeros(rp[-1], np.int32)
    for i in range(h):
        c = rng.choice(w, lens[i], replace=False)
        if sort:
            c.sort()
        ci[rp[i]:rp[i + 1]] = c
    v = (rng.random(rp[-1]) * 2 - 1).astype(dtype)
    return rp, ci, v


def main():
    R = pyoracle.Reference()
    os.makedirs(GOLD, exist_ok=True)
    mtx = os.path.join(GOLD, "ash85.mtx")
    shutil.copyfile(os.path.join(REF_DIR, "ash85.mtx"), mtx)  # data fixture
    h, w, rp, ci, v = R.read_mtx(mtx)
    out = dict(ash_h=h, ash_w=w, ash_rp=rp, ash_ci=ci, ash_v=v)
    meta = {"generator": "oracle/gen_golden.py", "source": "reference via oracle/ref_shim.cu",
            "ash85": {"h": h, "w": w, "nnz": int(rp[-1])}}

    # --- SpMV on ash85 (spmv.h:16-32)
    y = R.spmv(rp, ci, v, np.ones(w), np.ones(h), 1.0, 1.0)
    out["ash_spmv_ones"] = y
    meta["ash85"]["spmv_ones_sum"] = float(y.sum())
    xr = R.rand_fill(w)  # DenseVector(length) random ctor, matrix.h:659-665
    out["ash_x_rand"] = xr
    out["ash_spmv_rand_a3b4"] = R.spmv(rp, ci, v, xr, np.ones(h), 3.0, 4.0)

    # --- SpMM on ash85 (spmm.h:55-68), B = DenseMatrix(w,k,col_major) (matrix.h:519-529)
    for k, a, b in ((64, 1.0, 1.0), (256, 3.0, 4.0), (9, 0.5, 0.0)):
        B = R.rand_fill(w * k)
        Cm = R.spmm(rp, ci, v, k, B, np.ones(h * k), a, b)
        out["ash_spmm_k%d" % k] = Cm
        meta["ash85"]["spmm_k%d" % k] = {"alpha": a, "beta": b, "C0": float(Cm[0]),
                                         "B0": float(B[0])}

    # --- partitions (matrix.h:360-375) and dense splits (matrix.h:557-567)
    parts = {}
    for n in (1, 2, 3, 4, 8):
        p = R.partition_nnz(rp, n)
        parts[str(n)] = {"avg": p["avg"], "nnz_gpu": p["nnz_gpu"].tolist(),
                         "start_row": p["start_row"].tolist(), "stop_row": p["stop_row"].tolist()}
        for i, r in enumerate(p["rowptr"]):
            out["ash_part%d_rowptr%d" % (n, i)] = r
    meta["ash85"]["partition"] = parts
    meta["dense_split"] = {"%d/%d" % (k, n): [R.dense_split(k, n)[0], R.dense_split(k, n)[1].tolist()]
                           for k in (64, 256, 9, 85) for n in (1, 2, 4, 8)}

    # --- seeded synthetic matrices, both precisions
    rng = np.random.default_rng(211)
    cases = {
        "rnd_short": dict(h=300, w=300, avg=4, empty_frac=0.3),
        "rnd_mid": dict(h=257, w=257, avg=40, empty_frac=0.05),
        "rnd_long": dict(h=120, w=5000, avg=6, empty_frac=0.2, long_row=(57, 4100)),
    }
    for name, kw in cases.items():
        for dt, sfx in ((np.float64, "f64"), (np.float32, "f32")):
            hh, ww = kw["h"], kw["w"]
            a_rp, a_ci, a_v = rand_csr(rng, hh, ww, kw["avg"], kw.get("empty_frac", 0),
                                       kw.get("long_row"), dt)
            key = "%s_%s" % (name, sfx)
            out[key + "_rp"], out[key + "_ci"], out[key + "_v"] = a_rp, a_ci, a_v
            k = 12
            B = rng.random(ww * k).astype(dt)
            C0 = rng.random(hh * k).astype(dt)
            out[key + "_B"], out[key + "_C0"] = B, C0
            out[key + "_spmm"] = R.spmm(a_rp, a_ci, a_v, k, B, C0.copy(), dt(1.5), dt(-0.5))
            if hh == ww:  # the reference CPU SpMV asserts square A (spmv.h:21)
                x = rng.random(ww).astype(dt)
                y0 = rng.random(hh).astype(dt)
                out[key + "_x"], out[key + "_y0"] = x, y0
                out[key + "_spmv"] = R.spmv(a_rp, a_ci, a_v, x, y0.copy(), dt(1.5), dt(-0.5))
            p = R.partition_nnz(a_rp, 4)
            meta[key] = {"h": hh, "w": ww, "nnz": int(a_rp[-1]), "k": k, "alpha": 1.5,
                         "beta": -0.5, "part4": {"avg": p["avg"], "nnz_gpu": p["nnz_gpu"].tolist(),
                                                 "start_row": p["start_row"].tolist(),
                                                 "stop_row": p["stop_row"].tolist()}}

    np.savez_compressed(os.path.join(GOLD, "golden.npz"), **out)
    with open(os.path.join(GOLD, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print("wrote", GOLD, "(%d arrays)" % len(out))


if __name__ == "__main__":
    main()
