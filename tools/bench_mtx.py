#!/usr/bin/env python
"""Times the Matrix Market loader (CPU only): ours with 1..N host threads against the
reference's own fscanf loader (mmio_highlevel.h, via oracle/_ref when it was built) on a synthetic
unsorted `real general` file; checks that every run returns identical arrays.
    python tools/bench_mtx.py [--rows 1000000] [--nnz 8000000]"""
import argparse, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--nnz", type=int, default=8_000_000)
    a = ap.parse_args()
    import __graft_entry__ as ge
    pkg = ge.load_package()
    import pyoracle
    rng = np.random.default_rng(1)
    r, c = rng.integers(1, a.rows + 1, a.nnz), rng.integers(1, a.rows + 1, a.nnz)
    v = rng.standard_normal(a.nnz)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "big.mtx")
        with open(p, "w") as f:
            f.write("%%%%MatrixMarket matrix coordinate real general\n%d %d %d\n" % (a.rows, a.rows, a.nnz))
            np.savetxt(f, np.column_stack([r, c, v]), fmt="%d %d %.17g")
        print("file: %d rows, %d entries, %.1f MB; host cpus: %d" % (a.rows, a.nnz, os.path.getsize(p) / 1e6, os.cpu_count()))
        base = None
        for t in (1, 2, 4, 8, 16, 0):
            if t > (os.cpu_count() or 1):
                continue
            t0 = time.time()
            got = pkg.kernels.read_mtx(p, threads=t)
            dt = time.time() - t0
            print("ours, threads=%s: %.2f s" % (t or "auto", dt))
            if base is None:
                base = got
            else:
                assert all(np.array_equal(x, y) for x, y in zip(got[2:], base[2:]))
        if pyoracle.Reference.available():
            t0 = time.time()
            ref = pyoracle.Reference().read_mtx(p)
            print("reference loader (mmio_info + mmio_data, fscanf per entry): %.2f s" % (time.time() - t0))
            assert all(np.array_equal(x, y) for x, y in zip(ref[2:], base[2:]))
            print("arrays identical to the reference loader's: True")


if __name__ == "__main__":
    main()
