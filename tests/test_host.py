"""CPU-only tests of the product's host layer and C ABI (no compute calls):
 - the shared library loads and exports every symbol include/sblas_b200.h declares;
 - the integer partitioners / Matrix Market loader agree with the oracle and the goldens;
 - compute entry points fail loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol(sblas):
    hdr = open(os.path.join(ROOT, "include", "sblas_b200.h")).read()
    declared = set(re.findall(r"SBLAS_API\s+[\w\s\*]+?\b(sblas_\w+)\s*\(", hdr))
    assert len(declared) >= 38
    lib = sblas._lib.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), "libsblas_b200.so does not export " + name
    out = subprocess.check_output(["nm", "-D", "--defined-only", sblas._lib.LIB_PATH], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    extra = {s for s in exported if not s.startswith("sblas_")}
    assert not extra, "non-ABI symbols leak from the library: %s" % sorted(extra)[:5]


def test_abi_is_plain_c(sblas):
    """The boundary header must compile as C (no torch / C++ types in signatures)."""
    r = subprocess.run(["/usr/bin/gcc", "-std=c99", "-fsyntax-only", "-x", "c",
                        os.path.join(ROOT, "include", "sblas_b200.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_struct_layouts_match(sblas):
    src = r'''
    #include <stdio.h>
    #include <stddef.h>
    #include "sblas_b200.h"
    int main(){printf("%zu %zu %zu %zu %zu\n", sizeof(sblasOpts), offsetof(sblasOpts, replica),
      offsetof(sblasOpts, carry), sizeof(sblasCsrStats), offsetof(sblasCsrStats, spmm_tiles));return 0;}'''
    exe = "/tmp/sblas_layout_test"
    subprocess.run(["/usr/bin/gcc", "-I" + os.path.join(ROOT, "include"), "-x", "c", "-", "-o", exe],
                   input=src, text=True, check=True)
    got = list(map(int, subprocess.check_output([exe], text=True).split()))
    L = sblas._lib
    assert got == [C.sizeof(L.Opts), L.Opts.replica.offset, L.Opts.carry.offset,
                   C.sizeof(L.CsrStats), L.CsrStats.spmm_tiles.offset]


def test_mtx_loader_matches_reference_golden(sblas, golden):
    h, w, rp, ci, v = sblas.kernels.read_mtx(golden["ash85_path"])
    assert (h, w) == (85, 85)
    assert np.array_equal(rp, golden["ash_rp"])
    assert np.array_equal(ci, golden["ash_ci"])
    assert np.array_equal(v, golden["ash_v"])


def test_mtx_loader_kinds(sblas, oracle, tmp_path):
    cases = {
        "gen_real": "%%MatrixMarket matrix coordinate real general\n% c\n3 4 4\n1 1 1.5\n3 2 -2e0\n1 4 3\n2 2 .25\n",
        "sym_int": "%%MatrixMarket matrix coordinate integer symmetric\n3 3 4\n1 1 7\n3 1 2\n2 3 -4\n3 3 1\n",
        "herm_cplx": "%%MatrixMarket matrix coordinate complex hermitian\n2 2 2\n1 1 1.0 0.0\n2 1 2.5 -1.0\n",
        "skew": "%%MatrixMarket matrix coordinate real skew-symmetric\n3 3 2\n2 1 1.0\n3 2 2.0\n",
        "pat": "%%MatrixMarket matrix coordinate pattern general\n2 3 3\n2 3\n1 1\n2 1\n",
    }
    for name, txt in cases.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(txt)
        a = sblas.kernels.read_mtx(str(p))
        b = oracle.read_mtx(str(p))
        assert a[0] == b[0] and a[1] == b[1], name
        for x, y in zip(a[2:], b[2:]):
            assert np.array_equal(x, y), name


def test_mtx_loader_against_reference(sblas, reference, tmp_path):
    rng = np.random.default_rng(3)
    for sym in (False, True):
        n, nz = 40, 150
        r, c = rng.integers(1, n + 1, nz), rng.integers(1, n + 1, nz)
        if sym:
            r, c = np.maximum(r, c), np.minimum(r, c)
        p = tmp_path / ("r%d.mtx" % sym)
        with open(p, "w") as f:
            f.write("%%%%MatrixMarket matrix coordinate real %s\n%d %d %d\n" % (
                "symmetric" if sym else "general", n, n, nz))
            for i in range(nz):
                f.write("%d %d %.17g\n" % (r[i], c[i], rng.standard_normal()))
        a, b = sblas.kernels.read_mtx(str(p)), reference.read_mtx(str(p))
        for x, y in zip(a[2:], b[2:]):
            assert np.array_equal(x, y)


@pytest.mark.parametrize("kind", ["real general", "real symmetric", "pattern symmetric", "integer general",
                                  "complex hermitian"])
def test_mtx_loader_parallel_is_order_preserving(sblas, oracle, tmp_path, kind):
    """the multi-threaded parser + stable counting sort returns exactly what the serial
    loop (and the oracle's restatement of mmio_highlevel.h:130-281) returns, for any thread count, with
    unsorted entries, duplicates, comment lines and entries wrapped over lines in odd ways."""
    rng = np.random.default_rng(len(kind))
    field, symm = kind.split()
    n, nz = 300, 20000
    r, c = rng.integers(1, n + 1, nz), rng.integers(1, n + 1, nz)
    if symm != "general":
        r, c = np.maximum(r, c), np.minimum(r, c)
    seps = [" ", "  ", "\t", "\n", " \n ", "\r\n"]
    p = tmp_path / "m.mtx"
    with open(p, "w", newline="") as f:
        f.write("%%%%MatrixMarket matrix coordinate %s %s\n%% comment\n%%another\n %d %d %d\n" % (field, symm, n, n, nz))
        for i in range(nz):
            toks = [str(r[i]), str(c[i])]
            if field == "real":
                toks.append("%.17g" % rng.standard_normal())
            elif field == "integer":
                toks.append(str(int(rng.integers(-50, 50))))
            elif field == "complex":
                toks += ["%.6e" % rng.standard_normal(), "%g" % rng.standard_normal()]
            # mostly one entry per line, sometimes wrapped or run together like fscanf tolerates
            odd = rng.random() < 0.2
            f.write((seps[int(rng.integers(len(seps)))] if odd else " ").join(toks))
            f.write(seps[int(rng.integers(len(seps)))] if odd else "\n")
    ref = oracle.read_mtx(str(p))
    for t in (1, 2, 3, 5, 16):
        got = sblas.kernels.read_mtx(str(p), threads=t)
        assert got[0] == ref[0] and got[1] == ref[1]
        for x, y in zip(got[2:], ref[2:]):
            assert np.array_equal(x, y), (kind, t)
    sblas.kernels.read_mtx(str(p), threads=0)


def test_mtx_loader_size_line_forms(sblas, oracle, tmp_path):
    """The size line read as mm_read_mtx_crd_size reads it (mmio.h:338-366): an indented comment or a
    blank line before it, a fourth token on it (dropped with the line), and garbage instead of it."""
    head = "%%MatrixMarket matrix coordinate real general\n% c\n"
    body = "1 1 1.5\n3 2 -2.0\n"
    forms = {
        "indent": head + "  % indented comment\n3 3 2\n" + body,
        "blank": head + "\n3 3 2\n" + body,
        "fourth": head + "3 3 2 7\n" + body,
        "plain": head + "3 3 2\n" + body,
    }
    want = None
    for name, txt in forms.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(txt)
        got = sblas.kernels.read_mtx(str(p), threads=2)
        assert got[0] == 3 and got[1] == 3 and len(got[4]) == 2, name
        ref = oracle.read_mtx(str(p))
        for x, y in zip(got[2:], ref[2:]):
            assert np.array_equal(x, y), name
        if want is None:
            want = got
        for x, y in zip(got[2:], want[2:]):
            assert np.array_equal(x, y), name
    bad = tmp_path / "bad.mtx"
    bad.write_text(head + "three 3 2\n" + body)
    with pytest.raises(sblas.SblasError):
        sblas.kernels.read_mtx(str(bad))


def test_mtx_loader_size_line_forms_reference(sblas, reference, tmp_path):
    head = "%%MatrixMarket matrix coordinate real general\n"
    body = "1 1 1.5\n3 2 -2.0\n"
    for name, txt in {"indent": head + " % x\n3 3 2\n" + body, "fourth": head + "3 3 2 7\n" + body}.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(txt)
        a, b = sblas.kernels.read_mtx(str(p)), reference.read_mtx(str(p))
        assert a[0] == b[0] and a[1] == b[1], name
        for x, y in zip(a[2:], b[2:]):
            assert np.array_equal(x, y), name


def test_mtx_loader_short_file_is_an_error(sblas, tmp_path):
    p = tmp_path / "short.mtx"
    p.write_text("%%MatrixMarket matrix coordinate real general\n3 3 4\n1 1 1.0\n2 2 2.0\n3 3\n")
    with pytest.raises(sblas.SblasError):
        sblas.kernels.read_mtx(str(p), threads=2)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_partition_matches_golden(sblas, golden, n):
    p = sblas.kernels.csr_partition_nnz(golden["ash_rp"], n)
    g = golden["meta"]["ash85"]["partition"][str(n)]
    assert p["avg"] == g["avg"]
    assert p["nnz_gpu"].tolist() == g["nnz_gpu"]
    assert p["start_row"].tolist() == g["start_row"]
    assert p["stop_row"].tolist() == g["stop_row"]
    for i in range(n):
        assert np.array_equal(p["rowptr"][i], golden["ash_part%d_rowptr%d" % (n, i)])


def _rand_rowptr(rng, h, avg, empty=0.2, long_row=None):
    lens = rng.poisson(avg, h)
    lens[rng.random(h) < empty] = 0
    if long_row:
        lens[long_row[0]] = long_row[1]
    rp = np.zeros(h + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    return rp


def test_partition_matches_oracle_random(sblas, oracle):
    rng = np.random.default_rng(11)
    for trial in range(40):
        h = int(rng.integers(1, 400))
        rp = _rand_rowptr(rng, h, float(rng.uniform(0.3, 30)), float(rng.uniform(0, 0.6)),
                          (int(rng.integers(0, h)), int(rng.integers(0, 3000))) if trial % 3 == 0 else None)
        nnz = int(rp[-1])
        for n in (1, 2, 3, 4, 7, 8):
            if nnz < n:       # more GPUs than nnz: the reference produces -1 rows; tested below
                continue
            p = sblas.kernels.csr_partition_nnz(rp, n)
            q = oracle.partition_nnz(rp, n, float_ceil=False)
            assert p["avg"] == q["avg"]
            for f in ("nnz_gpu", "start_row", "stop_row"):
                assert np.array_equal(p[f], q[f]), (trial, n, f)
            for i in range(n):
                if p["nnz_gpu"][i] > 0:
                    assert np.array_equal(p["rowptr"][i], q["rowptr"][i])
        for idx in rng.integers(-2, nnz + 3, 30):
            assert sblas.kernels.csr_find_row(rp, int(idx)) == oracle.find_row(rp, int(idx))


def test_partition_degenerate_more_gpus_than_nnz(sblas, oracle):
    rp = np.array([0, 0, 2, 2, 3], np.int32)
    p = sblas.kernels.csr_partition_nnz(rp, 8)
    q = oracle.partition_nnz(rp, 8, float_ceil=False)   # literal reference arithmetic
    for f in ("nnz_gpu", "start_row", "stop_row"):
        assert np.array_equal(p[f], q[f])
    assert p["start_row"][5] == -1 and p["nnz_gpu"][4] <= 0


def test_dense_split_matches_oracle(sblas, oracle, golden):
    for key, (avg, dims) in golden["meta"]["dense_split"].items():
        k, n = map(int, key.split("/"))
        a, d = sblas.kernels.dense_split(k, n)
        assert a == avg and d.tolist() == dims
    for k in range(1, 70):
        for n in (1, 2, 3, 4, 8):
            a, d = sblas.kernels.dense_split(k, n)
            b, e = oracle.dense_split(k, n)
            assert a == b and np.array_equal(d, e)


def test_row_aligned_panels(sblas):
    rng = np.random.default_rng(5)
    for _ in range(20):
        h = int(rng.integers(1, 300))
        rp = _rand_rowptr(rng, h, 6.0, 0.3)
        for n in (1, 2, 4, 8):
            rb = sblas.kernels.csr_partition_rows(rp, n)
            assert rb[0] == 0 and rb[-1] == h and np.all(np.diff(rb) >= 0)
            avg = -(-int(rp[-1]) // n)
            for i in range(n):
                for r in range(rb[i], rb[i + 1]):        # every row starts inside its panel's window
                    assert i * avg <= rp[r] or i == 0
                    assert rp[r] < (i + 1) * avg or i == n - 1 or rp[r] == rp[-1]
    # exact split when rows are uniform: identical to the reference's nnz split
    rp = (np.arange(65, dtype=np.int32) * 32)
    assert sblas.kernels.csr_partition_rows(rp, 8).tolist() == list(range(0, 65, 8))
    p = sblas.kernels.csr_partition_nnz(rp, 8)
    assert p["start_row"].tolist() == list(range(0, 64, 8))


def test_compute_fails_loudly_without_gpu(sblas):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert sblas.kernels.device_count() == 0
    L = sblas._lib
    plan = C.c_void_p()
    rp = np.zeros(3, np.int32)
    rc = L.lib().sblas_plan_create_i32(None, C.c_int32(2), C.c_int64(0), rp.ctypes.data_as(C.c_void_p),
                                       C.byref(plan))
    assert rc != 0 and L.lib().sblas_last_error()          # error, not a silent CPU path
    with pytest.raises(L.SblasError):
        L.check(rc, "plan_create")
    # the entry points added in round 2 fail the same way (no silent host path): the allocation-free SpMV, the
    # streamed dense transfers, the managed allocator
    x = np.zeros(8)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = L.lib().sblas_spmv_csr_noalloc_f64(None, C.c_int32(2), C.c_int32(2), C.c_int64(0), vp(rp), vp(rp), vp(x), vp(x), vp(x),
                                            C.c_double(1.0), C.c_double(0.0))
    assert rc != 0 and L.lib().sblas_last_error()
    rc = L.lib().sblas_dense_upload_rows_f64(None, vp(x), C.c_int64(4), C.c_int64(0), C.c_int64(4), C.c_int32(2), vp(x), C.c_int64(2))
    assert rc != 0
    p = C.c_void_p()
    assert L.lib().sblas_managed_alloc(C.c_size_t(64), C.byref(p)) != 0
    # ... while the process-wide switches and the NVTX ranges are host-only and harmless without a device
    prev = L.lib().sblas_set_spmm_exact(C.c_int(1))
    assert L.lib().sblas_set_spmm_exact(C.c_int(prev)) == 1
    L.lib().sblas_range_push(b"test")
    L.lib().sblas_range_pop()


def test_python_mirror_api_surface(sblas):
    """Names of the reference's operator interface exist with the same arity."""
    import inspect
    for fn, n in ((sblas.sblas_spmv_csr_v1, 6), (sblas.sblas_spmm_csr_v1, 6), (sblas.sblas_spmm_csr_v2, 6)):
        assert len(inspect.signature(fn).parameters) == n
    assert (sblas.none, sblas.replicate, sblas.segment) == (0, 1, 2)       # matrix.h:34
    assert (sblas.row_major, sblas.col_major) == (0, 1)                    # matrix.h:37
    A = sblas.CsrSparseMatrix(os.path.join(ROOT, "tests", "golden", "ash85.mtx"))
    assert (A.height, A.width, A.nnz) == (85, 85, 523)
    for f in ("csrRowPtr", "csrColIdx", "csrVal", "csrRowPtr_gpu", "csrColIdx_gpu", "csrVal_gpu",
              "nnz_gpu", "starting_row_gpu", "stoping_row_gpu", "n_gpu", "policy"):
        assert hasattr(A, f)


def test_choose_stitch(sblas):
    """dist.choose_stitch: peer stores for even row panels, multicast for the skewed panels of a power-law input."""
    D = sblas.dist
    assert D.choose_stitch([0, 100, 200, 300, 400], 4) == "peer"
    assert D.choose_stitch([0, 10, 20], 2) == "peer"                      # two GPUs: nothing to replicate by switch
    c4_panels = [38929, 223252, 327772, 594328, 983560, 1527264, 1056772, 3636731]   # R-MAT config 4, 8 GPUs
    rb = np.concatenate([[0], np.cumsum(c4_panels)])
    assert D.choose_stitch(rb, 8) == "mc"


def test_mtx_loader_number_forms_bit_exact(sblas, oracle, tmp_path):
    """The loader's fast decimal path (<= 15 significant digits, |exponent| <= 22: one exactly rounded multiply or
    divide) and its strtod fallback return the same bits as the C library's parser, whatever the notation."""
    rng = np.random.default_rng(21)
    forms = ["%.3f", "%g", "%.10e", "%.15g", "%.16g", "%.17g", "%.1f", "%e", "%.12f", "%.15e"]
    toks = []
    for _ in range(4000):
        v = float(rng.standard_normal() * 10.0 ** int(rng.integers(-30, 30)))
        toks.append(forms[int(rng.integers(len(forms)))] % v)
    toks += ["0", "-0.0", "+1.5", ".5", "5.", "1e22", "1e23", "1e-22", "1e-23", "123456789012345", "1234567890123456",
             "999999999999999e22", "0.000001", "5e-324", "1.7976931348623157e308", "2.2250738585072014e-308",
             "000123.4500", "1E5", "1e+5", "12345678901234567890", "0.1", "0.2", "0.3", "1e0", "100000000000000000000000"]
    n = len(toks)
    p = tmp_path / "forms.mtx"
    with open(p, "w") as f:
        f.write("%%%%MatrixMarket matrix coordinate real general\n%d %d %d\n" % (n, 3, n))
        for i, t in enumerate(toks):
            f.write("%d %d %s\n" % (i + 1, 1 + i % 3, t))
    want = np.array([float(t) for t in toks])          # Python's float() is correctly rounded, like glibc strtod
    for t in (1, 4):
        h, w, rp, ci, v = sblas.kernels.read_mtx(str(p), threads=t)
        assert np.array_equal(v.view(np.uint64), want.view(np.uint64)), np.nonzero(v != want)[0][:5]
    ref = oracle.read_mtx(str(p))
    assert np.array_equal(ref[4].view(np.uint64), want.view(np.uint64))
