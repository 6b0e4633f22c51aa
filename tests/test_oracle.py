"""Pins the CPU oracle (oracle/sblas_oracle.c): against the committed golden vectors that
were generated from the reference's own code, against the known answers recorded in
BASELINE.md §2 / SURVEY.md §8(c), and — when oracle/_ref is built — against the reference
itself, bit for bit, on seeded random inputs."""
import numpy as np
import pytest


def test_mtx_reader_matches_golden(oracle, golden):
    h, w, rp, ci, v = oracle.read_mtx(golden["ash85_path"])
    assert (h, w, int(rp[-1])) == (85, 85, 523)          # mmio_highlevel.h:95-100 expansion
    assert np.array_equal(rp, golden["ash_rp"])
    assert np.array_equal(ci, golden["ash_ci"])           # file order, unsorted within row
    assert np.array_equal(v, golden["ash_v"])
    assert rp[:6].tolist() == [0, 5, 11, 15, 21, 25] and ci[:5].tolist() == [0, 1, 5, 6, 7]


def test_spmv_known_answers(oracle, golden):
    rp, ci, v = golden["ash_rp"], golden["ash_ci"], golden["ash_v"]
    y = oracle.spmv(rp, ci, v, np.ones(85), np.ones(85), 1.0, 1.0)
    assert np.array_equal(y, golden["ash_spmv_ones"])
    assert y[:4].tolist() == [6, 7, 5, 7] and y.sum() == 608.0
    assert np.array_equal(y, 1.0 + np.diff(rp))
    x = oracle.rand_fill(85)
    assert np.array_equal(x, golden["ash_x_rand"])        # glibc srand(211)/rand()
    y = oracle.spmv(rp, ci, v, x, np.ones(85), 3.0, 4.0)
    assert np.array_equal(y, golden["ash_spmv_rand_a3b4"])


@pytest.mark.parametrize("k,alpha,beta", [(64, 1.0, 1.0), (256, 3.0, 4.0), (9, 0.5, 0.0)])
def test_spmm_known_answers(oracle, golden, k, alpha, beta):
    rp, ci, v = golden["ash_rp"], golden["ash_ci"], golden["ash_v"]
    B = oracle.rand_fill(85 * k)
    C = oracle.spmm(rp, ci, v, k, B, np.ones(85 * k), alpha, beta)
    assert np.array_equal(C, golden["ash_spmm_k%d" % k])
    if k == 64:
        assert B[0] == 0.59819854358127278 and C[0] == 3.6699169691046314
        assert abs(C.sum() - 22129.266514203122) < 1e-8


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_partition_known_answers(oracle, golden, n):
    p = oracle.partition_nnz(golden["ash_rp"], n, float_ceil=True)
    q = oracle.partition_nnz(golden["ash_rp"], n, float_ceil=False)
    g = golden["meta"]["ash85"]["partition"][str(n)]
    for r in (p, q):
        assert r["avg"] == g["avg"]
        assert r["nnz_gpu"].tolist() == g["nnz_gpu"]
        assert r["start_row"].tolist() == g["start_row"]
        assert r["stop_row"].tolist() == g["stop_row"]
        for i in range(n):
            assert np.array_equal(r["rowptr"][i], golden["ash_part%d_rowptr%d" % (n, i)])
    if n == 2:   # row 39 straddles
        assert p["nnz_gpu"].tolist() == [262, 261] and p["start_row"].tolist() == [0, 39]
    if n == 8:
        assert p["nnz_gpu"].tolist() == [66] * 7 + [61]
        assert [len(x) for x in p["rowptr"]] == [12, 12, 12, 12, 12, 12, 14, 14]


def test_dense_split_known_answers(oracle, golden):
    for key, (avg, dims) in golden["meta"]["dense_split"].items():
        k, n = map(int, key.split("/"))
        a, d = oracle.dense_split(k, n)
        assert a == avg and d.tolist() == dims
    assert oracle.dense_split(64, 8)[1].tolist() == [8] * 8
    assert oracle.dense_split(9, 8)[1].tolist() == [2, 2, 2, 2, 1, -1, -3, -5]


@pytest.mark.parametrize("name", ["rnd_short", "rnd_mid", "rnd_long"])
@pytest.mark.parametrize("sfx", ["f64", "f32"])
def test_synthetic_goldens(oracle, golden, name, sfx):
    key = "%s_%s" % (name, sfx)
    m = golden["meta"][key]
    dt = np.float64 if sfx == "f64" else np.float32
    rp, ci, v = golden[key + "_rp"], golden[key + "_ci"], golden[key + "_v"]
    C = oracle.spmm(rp, ci, v, m["k"], golden[key + "_B"], golden[key + "_C0"].copy(),
                    dt(m["alpha"]), dt(m["beta"]))
    assert np.array_equal(C, golden[key + "_spmm"])
    if key + "_spmv" in golden:
        y = oracle.spmv(rp, ci, v, golden[key + "_x"], golden[key + "_y0"].copy(),
                        dt(m["alpha"]), dt(m["beta"]))
        assert np.array_equal(y, golden[key + "_spmv"])
    p = oracle.partition_nnz(rp, 4)
    assert p["nnz_gpu"].tolist() == m["part4"]["nnz_gpu"]
    assert p["start_row"].tolist() == m["part4"]["start_row"]
    assert p["stop_row"].tolist() == m["part4"]["stop_row"]


def test_omp_rows_bit_identical(oracle, golden):
    key = "rnd_mid_f64"
    rp, ci, v = golden[key + "_rp"], golden[key + "_ci"], golden[key + "_v"]
    h = len(rp) - 1
    y = oracle.spmv_rows_omp(0, h, rp, ci, v, golden[key + "_x"], golden[key + "_y0"].copy(),
                             1.5, -0.5, 4)
    assert np.array_equal(y, golden[key + "_spmv"])
    k = golden["meta"][key]["k"]
    C = oracle.spmm_rows_omp(0, h, rp, ci, v, k, golden[key + "_B"], golden["meta"][key]["w"],
                             golden[key + "_C0"].copy(), h, 1.5, -0.5, 4)
    assert np.array_equal(C, golden[key + "_spmm"])


def test_float_ceil_quirk(oracle):
    """the reference's float32 ceil drops trailing nnz when nnz > 2^24."""
    nnz, h = 134_217_733, 8
    rp = np.linspace(0, nnz, h + 1).astype(np.int32)
    p = oracle.partition_nnz(rp, 4, float_ceil=True)
    q = oracle.partition_nnz(rp, 4, float_ceil=False)
    assert int(p["nnz_gpu"].sum()) == nnz - 5          # reference loses 5 nnz
    assert int(q["nnz_gpu"].sum()) == nnz              # integer ceil-div (the product) does not


def test_against_reference_bitexact(oracle, reference):
    rng = np.random.default_rng(7)
    for dt in (np.float64, np.float32):
        for h, avg in ((64, 3), (200, 17), (33, 70)):
            lens = rng.poisson(avg, h)
            lens[rng.random(h) < 0.15] = 0
            lens = np.minimum(lens, h)
            rp = np.zeros(h + 1, np.int32)
            rp[1:] = np.cumsum(lens)
            ci = np.concatenate([rng.choice(h, l, replace=False) for l in lens] +
                                [np.zeros(0, np.int64)]).astype(np.int32)
            v = (rng.random(rp[-1]) * 2 - 1).astype(dt)
            x, y0 = rng.random(h).astype(dt), rng.random(h).astype(dt)
            a, b = dt(rng.random() * 3), dt(rng.random() * 3)
            assert np.array_equal(oracle.spmv(rp, ci, v, x, y0.copy(), a, b),
                                  reference.spmv(rp, ci, v, x, y0.copy(), a, b))
            k = 7
            B, C0 = rng.random(h * k).astype(dt), rng.random(h * k).astype(dt)
            assert np.array_equal(oracle.spmm(rp, ci, v, k, B, C0.copy(), a, b),
                                  reference.spmm(rp, ci, v, k, B, C0.copy(), a, b))
            if rp[-1] > 0:
                for n in (1, 2, 3, 5, 8):
                    if n > rp[-1]:
                        continue
                    p, q = oracle.partition_nnz(rp, n), reference.partition_nnz(rp, n)
                    assert p["avg"] == q["avg"]
                    for f in ("nnz_gpu", "start_row", "stop_row"):
                        assert np.array_equal(p[f], q[f])
                    for r1, r2 in zip(p["rowptr"], q["rowptr"]):
                        assert np.array_equal(r1, r2)
                for idx in rng.integers(0, rp[-1], 20):
                    assert oracle.find_row(rp, int(idx)) == reference.find_row(rp, int(idx))
    assert np.array_equal(oracle.rand_fill(1000), reference.rand_fill(1000))
    assert np.array_equal(oracle.rand_fill(1000, np.float32), reference.rand_fill(1000, np.float32))


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_format_conversion_against_reference(oracle, reference, dtype):
    """the restated CsrToCsc / CscToCsr (matrix.h:40-91) against the reference's own templates,
    with empty rows and columns, unsorted rows and duplicate (row, col) pairs."""
    rng = np.random.default_rng(5)
    for m, n, avg in ((1, 1, 1), (60, 45, 6), (200, 900, 3), (500, 20, 12)):
        lens = rng.poisson(avg, m) * (rng.random(m) > 0.2)
        rp = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        ci = rng.integers(0, n, rp[-1]).astype(np.int32)
        v = rng.standard_normal(rp[-1]).astype(dtype)
        a, b = oracle.csr2csc(m, n, rp, ci, v), reference.csr2csc(m, n, rp, ci, v)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
        if dtype == np.float64:
            a2, b2 = oracle.csc2csr(n, m, *a), reference.csc2csr(n, m, *b)
            assert all(np.array_equal(x, y) for x, y in zip(a2, b2))
            assert np.array_equal(a2[0], rp)


def test_coo_sort_restatement(oracle):
    rng = np.random.default_rng(6)
    r, c = rng.integers(0, 30, 2000).astype(np.int32), rng.integers(0, 30, 2000).astype(np.int32)
    v = rng.standard_normal(2000)
    rs, cs, vs = oracle.coo_sort(r, c, v)
    o = np.lexsort((c, r))   # stable: equal (row, col) pairs keep input order
    assert np.array_equal(rs, r[o]) and np.array_equal(cs, c[o]) and np.array_equal(vs, v[o])
