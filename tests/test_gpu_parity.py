"""Parity of the CUDA path (through the C ABI) against the CPU oracle / the reference goldens.

Bars (north star): bit-exact for integer work and for the sequential kernel paths; otherwise
per-row error <= 1e-12 (fp64) / 1e-5 (fp32) relative to (|A||x|)_i."""
import numpy as np
import pytest

import synth_ref as SR

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact_spmm(sblas):
    """These tests prove bit-identity with the reference's CPU loops: SpMM runs with separately rounded
    mul/add (sblas_set_spmm_exact).  The default fused multiply-add arithmetic is covered by the
    tolerance tests, which opt out with SBLAS_FLAG_FMA or switch the default back."""
    prev = sblas.kernels.set_spmm_exact(True)
    yield
    sblas.kernels.set_spmm_exact(prev)

TOL = {np.dtype(np.float64): 1e-12, np.dtype(np.float32): 1e-5}


def _t(dev, a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _rand_csr(rng, h, w, avg, empty=0.2, long_row=None, dtype=np.float64, sort=True):
    lens = rng.poisson(avg, h).astype(np.int64)
    lens[rng.random(h) < empty] = 0
    if long_row:
        lens[long_row[0]] = long_row[1]
    lens = np.minimum(lens, w)
    rp = np.zeros(h + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    ci = np.zeros(rp[-1], np.int32)
    for i in range(h):
        c = rng.choice(w, lens[i], replace=False)
        ci[rp[i]:rp[i + 1]] = np.sort(c) if sort else c
    v = (rng.random(rp[-1]) * 2 - 1).astype(dtype)
    return rp, ci, v


def _row_bound(oracle, rp, ci, v, x, alpha, beta, y0):
    """|alpha|*(|A||x|)_i + |beta*y0_i|: the scale the tolerance is relative to."""
    return abs(alpha) * oracle.absrow_spmv(rp, ci, v, x) + np.abs(beta * y0.astype(np.float64)) + 1e-300


def _check(got, ref, bound, dtype, exact=False):
    if exact:
        assert np.array_equal(got, ref), "not bit-identical: max |d| = %g" % np.abs(got - ref).max()
    else:
        err = np.abs(got.astype(np.float64) - ref.astype(np.float64)) / bound
        assert err.max() <= TOL[np.dtype(dtype)], "max rel row error %g" % err.max()


# ------------------------------------------------------------------------------------ SpMV ----
MATS = {
    "short": dict(h=3000, w=3000, avg=4, empty=0.3),
    "mid": dict(h=1500, w=1500, avg=40, empty=0.05),
    "long": dict(h=600, w=20000, avg=300, empty=0.0),
    "skew": dict(h=800, w=30000, avg=5, empty=0.4, long_row=(333, 25000)),
    "empty_tail": dict(h=2000, w=500, avg=0.2, empty=0.8),
    "tiny": dict(h=1, w=7, avg=3, empty=0.0),
}


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("mat", list(MATS))
@pytest.mark.parametrize("variant", ["auto", "tile", "pipe", "pipe2", "vector", "split", "splitd", "splits", "splitg", "splitg700", "splitgd", "splitm", "splitm700", "splitmd",
                                     "splitp", "noplan"])
def test_spmv_vs_oracle(sblas, dev, oracle, dtype, mat, variant):
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(hash((mat, variant)) % 1000)
    kw = MATS[mat]
    rp, ci, v = _rand_csr(rng, kw["h"], kw["w"], kw["avg"], kw["empty"], kw.get("long_row"), dtype)
    h, w = kw["h"], kw["w"]
    x = rng.random(w).astype(dtype)
    y0 = rng.random(h).astype(dtype)
    alpha, beta = dtype(1.25), dtype(-0.75)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), alpha, beta)
    drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
    plan = None if variant == "noplan" else K.Plan(drp, len(ci))
    # split = sub-warp per row phase 2 (default), splitd = its default tile, splitm* = balanced merge-path phase 2
    # (tile 512, a tile that is not a multiple of the CTA size, the default tile), splits = row pointers before the
    # A stream (first version), splitp = staged flavour
    var = dict(auto=0, noplan=0, tile=sblas.SPMV_TILE, vector=sblas.SPMV_VECTOR, split=sblas.SPMV_SPLIT,
               splitd=sblas.SPMV_SPLIT, splits=sblas.SPMV_SPLIT, splitg=sblas.SPMV_SPLIT, splitg700=sblas.SPMV_SPLIT,
               splitgd=sblas.SPMV_SPLIT, splitm=sblas.SPMV_SPLIT, splitm700=sblas.SPMV_SPLIT, splitmd=sblas.SPMV_SPLIT,
               splitp=sblas.SPMV_SPLIT, pipe=sblas.SPMV_PIPE, pipe2=sblas.SPMV_PIPE)[variant]
    opts = sblas.make_opts(variant=var, tile={"tile": 512, "split": 512, "splits": 512, "splitg": 512, "splitg700": 700, "splitm": 512, "splitm700": 700, "splitp": 512,
                                              "pipe": 320, "pipe2": 3072}.get(variant, 0),
                           stages={"pipe": 4, "pipe2": 2, "splitp": 2}.get(variant, 0),
                           flags=sblas.FLAG_SPLIT_MERGE if variant.startswith("splitm") else
                           sblas.FLAG_SPLIT_SEG if variant.startswith("splitg") else
                           sblas.FLAG_SPLIT_SERIAL if variant == "splits" else 0)
    K.spmv(drp, dci, dv, dx, dy, alpha, beta, plan=plan, opts=opts)
    torch.cuda.synchronize()
    bound = _row_bound(oracle, rp, ci, v, x, alpha, beta, y0)
    _check(dy.cpu().numpy(), ref, bound, dtype)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_spmv_short_rows_bit_identical(sblas, dev, oracle, dtype):
    """Rows reduced by one lane in CSR order with rounded mul/add == the reference CPU loop."""
    import torch
    K = sblas.kernels
    rp, ci, v = SR.laplace2d(97, 53, dtype)
    n = 97 * 53
    x = SR.fill_uniform(n, 211, 7, 0, 1, dtype)
    for alpha, beta in ((1.0, 1.0), (3.0, 4.0), (0.5, 0.0)):
        ref = oracle.spmv(rp, ci, v, x, np.ones(n, dtype), dtype(alpha), dtype(beta))
        drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, np.ones(n, dtype)))
        K.spmv(drp, dci, dv, dx, dy, alpha, beta, plan=K.Plan(drp, len(ci)))
        torch.cuda.synchronize()
        assert np.array_equal(dy.cpu().numpy(), ref)
        dy2 = _t(dev, np.ones(n, dtype))
        K.spmv(drp, dci, dv, dx, dy2, alpha, beta, plan=K.Plan(drp, len(ci)),
               opts=sblas.make_opts(variant=sblas.SPMV_TILE, lanes=1, tile=777))
        assert np.array_equal(dy2.cpu().numpy(), ref)


def test_spmv_golden_ash85(sblas, dev, golden):
    """BASELINE config 1: ash85, x = y = 1, alpha = beta = 1 -> y = 1 + rowlen, sum 608."""
    import torch
    K = sblas.kernels
    h, w, rp, ci, v = K.read_mtx(golden["ash85_path"])
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    dy = torch.ones(h, dtype=torch.float64, device=dev)
    K.spmv(drp, dci, dv, torch.ones(w, dtype=torch.float64, device=dev), dy, 1.0, 1.0,
           plan=K.Plan(drp, len(ci)))
    y = dy.cpu().numpy()
    assert np.array_equal(y, golden["ash_spmv_ones"]) and y.sum() == 608.0
    dy = torch.ones(h, dtype=torch.float64, device=dev)
    K.spmv(drp, dci, dv, _t(dev, golden["ash_x_rand"]), dy, 3.0, 4.0, plan=K.Plan(drp, len(ci)))
    assert np.array_equal(dy.cpu().numpy(), golden["ash_spmv_rand_a3b4"])


@pytest.mark.parametrize("name", ["rnd_short", "rnd_mid"])
@pytest.mark.parametrize("sfx", ["f64", "f32"])
def test_spmv_reference_goldens(sblas, dev, oracle, golden, name, sfx):
    K = sblas.kernels
    key = "%s_%s" % (name, sfx)
    m = golden["meta"][key]
    dt = np.float64 if sfx == "f64" else np.float32
    rp, ci, v, x, y0 = (golden[key + s] for s in ("_rp", "_ci", "_v", "_x", "_y0"))
    drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
    K.spmv(drp, dci, dv, dx, dy, m["alpha"], m["beta"], plan=K.Plan(drp, len(ci)))
    bound = _row_bound(oracle, rp, ci, v, x, m["alpha"], m["beta"], y0)
    _check(dy.cpu().numpy(), golden[key + "_spmv"], bound, dt)


def test_spmv_partials_and_replicas(sblas, dev, oracle):
    """HEAD/TAIL partial rows go to the carry buffer, everything else to y and every replica."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(5)
    rp, ci, v = _rand_csr(rng, 500, 500, 9, 0.1)
    x, y0 = rng.random(500), rng.random(500)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), 2.0, 0.5)
    raw = oracle.spmv(rp, ci, v, x, np.zeros(500), 1.0, 0.0)
    for var in (sblas.SPMV_TILE, sblas.SPMV_VECTOR, sblas.SPMV_SPLIT, sblas.SPMV_PIPE):
        drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
        rep = torch.full((500,), -7.0, dtype=torch.float64, device=dev)
        carry = torch.zeros(8, dtype=torch.float64, device=dev)
        o = sblas.make_opts(variant=var, flags=sblas.FLAG_HEAD_PARTIAL | sblas.FLAG_TAIL_PARTIAL,
                            replicas=[rep.data_ptr()], carry=carry.data_ptr(), tile=256)
        K.spmv(drp, dci, dv, dx, dy, 2.0, 0.5, plan=K.Plan(drp, len(ci)), opts=o)
        y, r, c = dy.cpu().numpy(), rep.cpu().numpy(), carry.cpu().numpy()
        bound = _row_bound(oracle, rp, ci, v, x, 2.0, 0.5, y0)
        _check(y[1:-1], ref[1:-1], bound[1:-1], np.float64)
        assert np.array_equal(r[1:-1], y[1:-1])
        assert y[0] == y0[0] and y[-1] == y0[-1] and r[0] == -7.0 and r[-1] == -7.0
        assert abs(c[0] - raw[0]) <= 1e-12 * bound[0] and abs(c[4] - raw[-1]) <= 1e-12 * bound[-1]


# ------------------------------------------------------------------------------------ SpMM ----
def _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, alpha, beta):
    return oracle.spmm(rp, ci, v, k, Bc, C0c.copy(), alpha, beta)


def _to_rm(a_cm, rows, k, ld):
    out = np.zeros((rows, ld), a_cm.dtype)
    out[:, :k] = a_cm.reshape(k, rows).T
    return out.reshape(-1)


def _from_rm(a_rm, rows, k, ld):
    return np.ascontiguousarray(a_rm.reshape(rows, ld)[:, :k].T).reshape(-1)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("k", [1, 3, 8, 12, 32, 64, 100, 256, 700])
@pytest.mark.parametrize("mat", ["short", "mid", "skew"])
def test_spmm_rowmajor_bit_identical(sblas, dev, oracle, dtype, k, mat):
    """Every (row, column) sum is accumulated by one lane in CSR order with rounded mul/add:
    bit-identical to sblas_spmm_csr_cpu (spmm.h:55-68)."""
    import torch
    K = sblas.kernels
    kw = dict(MATS[mat])
    if mat == "skew":
        kw.update(h=200, w=3000, long_row=(77, 2500))
    else:
        kw.update(h=min(kw["h"], 400), w=min(kw["w"], 400))
    rng = np.random.default_rng(k)
    rp, ci, v = _rand_csr(rng, kw["h"], kw["w"], kw["avg"], kw["empty"], kw.get("long_row"), dtype,
                          sort=False)
    h, w = kw["h"], kw["w"]
    Bc, C0c = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    alpha, beta = dtype(1.5), dtype(-0.5)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, alpha, beta)
    vec = 16 // np.dtype(dtype).itemsize
    for ld in ((k + vec - 1) // vec * vec, k + 1 if k % vec == 0 else k):   # vector path, scalar path
        dB, dC = _t(dev, _to_rm(Bc, w, k, ld)), _t(dev, _to_rm(C0c, h, k, ld))
        drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
        K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC, ld, alpha, beta, w, plan=K.Plan(drp, len(ci)))
        torch.cuda.synchronize()
        got = _from_rm(dC.cpu().numpy(), h, k, ld)
        assert np.array_equal(got, ref), "k=%d ld=%d max|d|=%g" % (k, ld, np.abs(got - ref).max())


def _banded_csr(rng, n, bw, dtype, jitter=0):
    """row i: bw consecutive columns starting at clamp(i - bw/2), optionally with random holes"""
    c0 = np.clip(np.arange(n) - bw // 2, 0, n - bw)
    cols = (c0[:, None] + np.arange(bw)[None, :])
    keep = rng.random(cols.shape) >= jitter
    keep[:, 0] = True
    lens = keep.sum(1)
    rp = np.zeros(n + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    ci = cols[keep].astype(np.int32)
    v = (rng.random(len(ci)) * 2 - 1).astype(dtype)
    return rp, ci, v


@pytest.mark.parametrize("dtype,k", [(np.float64, 64), (np.float64, 256), (np.float64, 100), (np.float64, 300),
                                     (np.float64, 40), (np.float32, 128), (np.float32, 200), (np.float32, 72)])
@pytest.mark.parametrize("mat", ["banded", "banded_holes", "unsorted", "ragged_tail"])
@pytest.mark.parametrize("piped", [False, True, "tileb"])
@pytest.mark.parametrize("R", [4, 8])
def test_spmm_rowblock_bit_identical(sblas, dev, oracle, dtype, k, mat, piped, R):
    """SBLAS_SPMM_ROWBLOCK: the R x 1 row-block mirror of A (one B load feeds up to R adjacent rows)
    accumulates every row's nonzeros in CSR storage order with rounded mul/add, so it is
    bit-identical to sblas_spmm_csr_cpu (spmm.h:55-68) — also for rows whose columns are unsorted."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(k + len(mat))
    if mat == "banded":
        h = w = 517
        rp, ci, v = _banded_csr(rng, h, 24, dtype)
    elif mat == "banded_holes":
        h = w = 400
        rp, ci, v = _banded_csr(rng, h, 16, dtype, jitter=0.3)
    elif mat == "unsorted":          # .mtx file order (mmio_highlevel.h keeps it): shuffle within rows
        h = w = 300
        rp, ci, v = _banded_csr(rng, h, 12, dtype)
        for r in range(h):
            perm = rng.permutation(rp[r + 1] - rp[r]) + rp[r]
            ci[rp[r]:rp[r + 1]], v[rp[r]:rp[r + 1]] = ci[perm], v[perm]
    else:                            # rows % 4 != 0, empty rows, a dense-ish stretch
        h, w = 203, 150
        rp, ci, v = _rand_csr(rng, h, w, 40, 0.2, None, dtype)
    Bc, C0c = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    alpha, beta = dtype(1.5), dtype(-0.5)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, alpha, beta)
    vec = 16 // np.dtype(dtype).itemsize
    ld = (k + vec - 1) // vec * vec
    dB, dC = _t(dev, _to_rm(Bc, w, k, ld)), _t(dev, _to_rm(C0c, h, k, ld))
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    plan = K.Plan(drp, len(ci))
    Rr, nb, fill = plan.build_rowblocks(dci, dv, R=R)
    if mat == "banded":
        assert Rr == R and nb > 0 and (0.7 if R == 8 else 0.5) <= fill <= 1.0
    if nb == 0:  # low fill: the mirror is normally dropped (CSR kernel); force it to test the arithmetic
        Rr, nb, fill = plan.build_rowblocks(dci, dv, R=-R)
    assert Rr == R and nb > 0 and 1.0 / R <= fill <= 1.0
    var = {False: sblas.SPMM_ROWBLOCK, True: sblas.SPMM_ROWBLOCK_PIPE, "tileb": sblas.SPMM_ROWBLOCK_TILEB}[piped]
    if piped == "tileb" and (R != 4 or mat in ("unsorted", "ragged_tail")):
        # B rows staged per tile need the mirror's tile column lists: R = 4, ascending columns within
        # rows, and a tile's list must fit shared memory; otherwise the call fails loudly
        if R != 4 or mat == "unsorted":
            with pytest.raises(sblas.SblasError, match="TILEB"):
                K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC, ld, alpha, beta, w, plan=plan,
                       opts=sblas.make_opts(variant=var))
            return
    try:
        K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC, ld, alpha, beta, w, plan=plan,
               opts=sblas.make_opts(variant=var, stages=3 if k == 100 else 0))
    except sblas.SblasError as e:
        if piped == "tileb" and mat == "ragged_tail" and "TILEB" in str(e):
            pytest.skip("tile column list does not fit shared memory at this k")
        raise
    torch.cuda.synchronize()
    got = _from_rm(dC.cpu().numpy(), h, k, ld)
    assert np.array_equal(got, ref), "k=%d max|d|=%g" % (k, np.abs(got - ref).max())
    if piped is True:   # column slices per CTA forced to 1 / 2 / 4 (tile of 8 / 4 / 2 block rows)
        for P in (1, 2, 4):
            dCp = _t(dev, _to_rm(C0c, h, k, ld))
            K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dCp, ld, alpha, beta, w, plan=plan,
                   opts=sblas.make_opts(variant=var, lanes=P,
                                        flags=sblas.FLAG_PREFETCH_B if P == 2 else sblas.FLAG_NO_PREFETCH_B))
            assert np.array_equal(_from_rm(dCp.cpu().numpy(), h, k, ld), ref), "P=%d" % P
    # AUTO takes the same path when the plan holds the mirror; FMA flag stays within tolerance
    dC2 = _t(dev, _to_rm(C0c, h, k, ld))
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC2, ld, alpha, beta, w, plan=plan)
    assert np.array_equal(_from_rm(dC2.cpu().numpy(), h, k, ld), ref)
    dC3 = _t(dev, _to_rm(C0c, h, k, ld))
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC3, ld, alpha, beta, w, plan=plan,
           opts=sblas.make_opts(variant=var, flags=sblas.FLAG_FMA))
    got3 = _from_rm(dC3.cpu().numpy(), h, k, ld)
    tol = 1e-12 if dtype == np.float64 else 1e-5
    for n in range(0, k, max(1, k // 7)):
        b = _row_bound(oracle, rp, ci, v, Bc[n * w:(n + 1) * w], alpha, beta, C0c[n * h:(n + 1) * h])
        assert np.all(np.abs(got3[n * h:(n + 1) * h] - ref[n * h:(n + 1) * h]) <= tol * b)


def test_rowblock_height_selection(sblas, dev):
    """R = 0 (what sync2gpu asks for): the R = 4 mirror when its fill pays, else no mirror."""
    K = sblas.kernels
    rng = np.random.default_rng(5)
    for kind, want in (("band", 4), ("holes", 4), ("random", 0)):
        if kind == "band":
            rp, ci, v = _banded_csr(rng, 600, 24, np.float64)
        elif kind == "holes":
            rp, ci, v = _banded_csr(rng, 600, 6, np.float64, jitter=0.25)
        else:
            rp, ci, v = _rand_csr(rng, 600, 600, 8, 0.0, None, np.float64)
        drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
        plan = K.Plan(drp, len(ci))
        R, nb, fill = plan.build_rowblocks(dci, dv)
        assert (R if nb else 0) == want, (kind, R, nb, fill)


@pytest.mark.parametrize("piped", [False, True, "bulk"])
def test_spmm_rowblock_partials_and_replicas(sblas, dev, oracle, piped):
    """HEAD/TAIL partial rows go to the carry buffer, the rest to C and every replica ("bulk": the replica is written
    by TMA bulk stores of the row slices, SBLAS_FLAG_STITCH_BULK)."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(11)
    h = w = 130
    k = 64
    rp, ci, v = _banded_csr(rng, h, 10, np.float64)
    Bc, C0c = rng.random(w * k), rng.random(h * k)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, 2.0, 0.5)
    raw = _spmm_ref(oracle, rp, ci, v, k, Bc, np.zeros(h * k), 1.0, 0.0)
    dB, dC = _t(dev, _to_rm(Bc, w, k, k)), _t(dev, _to_rm(C0c, h, k, k))
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    rep = torch.full((h * k,), -7.0, dtype=torch.float64, device=dev)
    carry = torch.zeros(2 * k, dtype=torch.float64, device=dev)
    plan = K.Plan(drp, len(ci))
    assert plan.build_rowblocks(dci, dv)[1] > 0
    o = sblas.make_opts(variant=sblas.SPMM_ROWBLOCK_PIPE if piped else sblas.SPMM_ROWBLOCK,
                        flags=sblas.FLAG_HEAD_PARTIAL | sblas.FLAG_TAIL_PARTIAL | (sblas.FLAG_STITCH_BULK if piped == "bulk" else 0),
                        replicas=[rep.data_ptr()], carry=carry.data_ptr())
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, k, dC, k, 2.0, 0.5, w, plan=plan, opts=o)
    torch.cuda.synchronize()
    got = _from_rm(dC.cpu().numpy(), h, k, k).reshape(k, h)
    r = _from_rm(rep.cpu().numpy(), h, k, k).reshape(k, h)
    refm, rawm, c0m = ref.reshape(k, h), raw.reshape(k, h), C0c.reshape(k, h)
    assert np.array_equal(got[:, 1:-1], refm[:, 1:-1]) and np.array_equal(r[:, 1:-1], refm[:, 1:-1])
    assert np.array_equal(got[:, 0], c0m[:, 0]) and np.array_equal(got[:, -1], c0m[:, -1])
    assert np.all(r[:, 0] == -7.0) and np.all(r[:, -1] == -7.0)
    c = carry.cpu().numpy()
    assert np.array_equal(c[:k], rawm[:, 0]) and np.array_equal(c[k:], rawm[:, -1])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("k", [1, 5, 64])
@pytest.mark.parametrize("mat", ["short", "mid"])
def test_spmm_colmajor_vs_oracle(sblas, dev, oracle, dtype, k, mat):
    import torch
    K = sblas.kernels
    kw = MATS[mat]
    h = w = 500
    rng = np.random.default_rng(k + 17)
    rp, ci, v = _rand_csr(rng, h, w, kw["avg"], kw["empty"], None, dtype)
    Bc, C0c = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    alpha, beta = dtype(0.75), dtype(2.0)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, alpha, beta)
    drp, dci, dv, dB, dC = (_t(dev, a) for a in (rp, ci, v, Bc, C0c))
    K.spmm(drp, dci, dv, k, sblas.COL_MAJOR, dB, w, dC, h, alpha, beta, w, plan=K.Plan(drp, len(ci)))
    torch.cuda.synchronize()
    got = dC.cpu().numpy()
    # bound per (row, column): reuse the SpMV bound column by column
    for n in range(k):
        b = _row_bound(oracle, rp, ci, v, Bc[n * w:(n + 1) * w], alpha, beta, C0c[n * h:(n + 1) * h])
        _check(got[n * h:(n + 1) * h], ref[n * h:(n + 1) * h], b, dtype)


def test_spmm_fma_flag_within_tolerance(sblas, dev, oracle):
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(2)
    h = w = 300
    k = 64
    rp, ci, v = _rand_csr(rng, h, w, 20, 0.1)
    Bc, C0c = rng.random(w * k), rng.random(h * k)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, 1.0, 1.0)
    dB, dC = _t(dev, _to_rm(Bc, w, k, k)), _t(dev, _to_rm(C0c, h, k, k))
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, k, dC, k, 1.0, 1.0, w, plan=K.Plan(drp, len(ci)),
           opts=sblas.make_opts(flags=sblas.FLAG_FMA))
    got = _from_rm(dC.cpu().numpy(), h, k, k)
    for n in range(0, k, 9):
        b = _row_bound(oracle, rp, ci, v, Bc[n * w:(n + 1) * w], 1.0, 1.0, C0c[n * h:(n + 1) * h])
        _check(got[n * h:(n + 1) * h], ref[n * h:(n + 1) * h], b, np.float64)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("k", [64, 256])
@pytest.mark.parametrize("variant", ["auto_csr", "rowmajor", "colmajor", "rowblock", "rowblock_pipe", "rowblock_tileb"])
def test_spmm_default_arithmetic_within_tolerance(sblas, dev, oracle, dtype, k, variant):
    """The DEFAULT arithmetic of SpMM is fused multiply-add: every kernel variant, k = 64
    and 256, stays within the north star's per-row bound (fp64 1e-12, fp32 1e-5, relative to |A||B|), and the
    default really is FMA (it differs from the exact result somewhere, and equals the SBLAS_FLAG_FMA run)."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(k + len(variant))
    h = w = 640
    rp, ci, v = _banded_csr(rng, h, 24, dtype, jitter=0.1)
    Bc, C0c = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    alpha, beta = dtype(1.25), dtype(-0.75)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, alpha, beta)
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    plan = K.Plan(drp, len(ci))
    var = {"auto_csr": 0, "rowmajor": sblas.SPMM_ROWMAJOR, "colmajor": sblas.SPMM_COLMAJOR, "rowblock": sblas.SPMM_ROWBLOCK,
           "rowblock_pipe": sblas.SPMM_ROWBLOCK_PIPE, "rowblock_tileb": sblas.SPMM_ROWBLOCK_TILEB}[variant]
    if variant.startswith("rowblock"):
        plan.build_rowblocks(dci, dv, R=-4)
    cm = variant == "colmajor"

    def run(flags):
        if cm:
            dB, dC = _t(dev, Bc), _t(dev, C0c)
            K.spmm(drp, dci, dv, k, sblas.COL_MAJOR, dB, w, dC, h, alpha, beta, w, plan=plan,
                   opts=sblas.make_opts(variant=var, flags=flags))
            return dC.cpu().numpy()
        dB, dC = _t(dev, _to_rm(Bc, w, k, k)), _t(dev, _to_rm(C0c, h, k, k))
        K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, k, dC, k, alpha, beta, w, plan=plan,
               opts=sblas.make_opts(variant=var, flags=flags))
        return _from_rm(dC.cpu().numpy(), h, k, k)

    prev = K.set_spmm_exact(False)          # the library's initial default
    try:
        got_default = run(0)
    finally:
        K.set_spmm_exact(prev)
    got_fma, got_exact = run(sblas.FLAG_FMA), run(sblas.FLAG_EXACT)
    torch.cuda.synchronize()
    assert np.array_equal(got_default, got_fma)
    if not cm:      # one lane per (row, column) in CSR order; the col-major kernel shuffles-reduces lanes
        assert np.array_equal(got_exact, ref)
    assert not np.array_equal(got_default, got_exact), "default arithmetic is not FMA"
    for n in range(0, k, max(1, k // 9)):
        b = _row_bound(oracle, rp, ci, v, Bc[n * w:(n + 1) * w], alpha, beta, C0c[n * h:(n + 1) * h])
        _check(got_default[n * h:(n + 1) * h], ref[n * h:(n + 1) * h], b, dtype)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("k", [1, 3, 5, 9, 66])
@pytest.mark.parametrize("rowblock", [False, True])
def test_spmm_leaves_padding_columns_alone(sblas, dev, oracle, dtype, k, rowblock):
    """ADVICE r1 (spmm.cu vector path): with k not a multiple of the 16-byte vector and ld >= roundup(k), the
    columns [k, ld) of C (a neighbouring sub-matrix, or padding) must not be touched — in C or in a replica."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(k)
    h, w = 333, 257
    rp, ci, v = _rand_csr(rng, h, w, 9, 0.1, None, dtype)
    vec = 16 // np.dtype(dtype).itemsize
    ld = (k + vec - 1) // vec * vec + vec          # a partial vector AND a whole untouched one
    Bc, C0c = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    ref = _spmm_ref(oracle, rp, ci, v, k, Bc, C0c, dtype(2.0), dtype(0.5))
    B = rng.random((w, ld)).astype(dtype)          # garbage (not zeros) behind column k of B too
    B[:, :k] = Bc.reshape(k, w).T
    C0 = np.full((h, ld), -7.0, dtype)
    C0[:, :k] = C0c.reshape(k, h).T
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    plan = K.Plan(drp, len(ci))
    if rowblock:
        plan.build_rowblocks(dci, dv, R=-4)
    dB, dC, dR = _t(dev, B.reshape(-1)), _t(dev, C0.reshape(-1)), _t(dev, np.full(h * ld, -9.0, dtype))
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC, ld, 2.0, 0.5, w, plan=plan,
           opts=sblas.make_opts(variant=sblas.SPMM_ROWBLOCK if rowblock and k >= 16 * vec else 0, replicas=[dR.data_ptr()]))
    torch.cuda.synchronize()
    got, rep = dC.cpu().numpy().reshape(h, ld), dR.cpu().numpy().reshape(h, ld)
    assert np.array_equal(np.ascontiguousarray(got[:, :k].T).reshape(-1), ref)
    assert np.all(got[:, k:] == -7.0), "columns >= k of C were overwritten"
    assert np.array_equal(rep[:, :k], got[:, :k]) and np.all(rep[:, k:] == -9.0), "columns >= k of the replica were overwritten"


@pytest.mark.parametrize("k", [64, 256, 9])
def test_spmm_golden_ash85(sblas, dev, golden, oracle, k):
    """BASELINE config 1 (k=64, alpha=beta=1) and unit_test.cu:183 (k=256, alpha=3, beta=4)."""
    K = sblas.kernels
    m = golden["meta"]["ash85"]["spmm_k%d" % k]
    h, w, rp, ci, v = K.read_mtx(golden["ash85_path"])
    Bc = oracle.rand_fill(w * k)            # DenseMatrix(w, k, col_major) ctor, seed 211
    ld = (k + 1) // 2 * 2
    dB, dC = _t(dev, _to_rm(Bc, w, k, ld)), _t(dev, _to_rm(np.ones(h * k), h, k, ld))
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    K.spmm(drp, dci, dv, k, sblas.ROW_MAJOR, dB, ld, dC, ld, m["alpha"], m["beta"], w,
           plan=K.Plan(drp, len(ci)))
    got = _from_rm(dC.cpu().numpy(), h, k, ld)
    assert np.array_equal(got, golden["ash_spmm_k%d" % k])
    if k == 64:
        assert got[0] == 3.6699169691046314


# ---------------------------------------------------------------------------- utility kernels ----
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_axpby_scal_transpose(sblas, dev, oracle, dtype):
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(1)
    for n in (1, 7, 1000, 100003):
        a, b = rng.random(n).astype(dtype), rng.random(n).astype(dtype)
        ref = oracle.axpby(a.copy(), b, dtype(3.0), dtype(4.0))
        da = _t(dev, a)
        K.axpby(da, _t(dev, b), 3.0, 4.0)
        assert np.array_equal(da.cpu().numpy(), ref)                      # kernel.h:34 order
        db = _t(dev, a)
        K.scal_add(db, 0.5, 2.0)
        assert np.array_equal(db.cpu().numpy(), a * dtype(2.0) + dtype(0.5))
    for (r, c, ld_in, ld_out) in ((5, 3, 3, 5), (100, 64, 64, 100), (33, 65, 70, 40), (1000, 7, 7, 1003)):
        src = rng.random(r * ld_in).astype(dtype)
        dst = torch.zeros(c * ld_out, dtype=_t(dev, src).dtype, device=dev)
        K.transpose(_t(dev, src), ld_in, r, c, dst, ld_out)
        got = dst.cpu().numpy().reshape(c, ld_out)[:, :r]
        assert np.array_equal(got, src.reshape(r, ld_in)[:, :c].T)


# ------------------------------------------------------------------------------- generators ----
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_synth_generators_match_numpy(sblas, dev, dtype):
    import torch
    S = sblas.synth
    td = torch.float64 if dtype == np.float64 else torch.float32
    t = S.fill_uniform(torch.empty(1000, dtype=td, device=dev), 211, 5, -1.0, 1.0)
    assert np.array_equal(t.cpu().numpy(), SR.fill_uniform(1000, 211, 5, -1.0, 1.0, dtype))
    for got, ref in ((S.laplace2d(37, 29, td, dev), SR.laplace2d(37, 29, dtype)),
                     (S.uniform(300, 5000, 32, td, dev), SR.uniform(300, 5000, 32, 211, dtype)),
                     (S.banded(500, 64, td, dev), SR.banded(500, 64, 211, dtype)),
                     (S.rmat(10, 8, td, dev), SR.rmat(10, 8, 211, dtype))):
        for g, r in zip(got[2:], ref):
            assert np.array_equal(g.cpu().numpy(), r)
    assert S.laplace2d(2048, 2048, td, dev)[4].numel() == 20963328


# ------------------------------------------------------ format containers ----
@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1, 1), (300, 200, 5), (50, 40000, 30), (20000, 64, 3), (4000, 4000, 0.01)])
def test_csr2csc_bit_identical(sblas, dev, oracle, dtype, shape):
    """CsrToCsc / CscToCsr on the device == the reference's CPU loops (matrix.h:40-91), bit for bit: empty rows
    and columns, unsorted rows, duplicate (row, col) pairs."""
    import torch
    K = sblas.kernels
    m, n, avg = shape
    rng = np.random.default_rng(m + n)
    lens = rng.poisson(avg, m) * (rng.random(m) > 0.2)
    rp = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    ci = rng.integers(0, n, rp[-1]).astype(np.int32)
    v = rng.standard_normal(rp[-1]).astype(dtype)
    ref = oracle.csr2csc(m, n, rp, ci, v)
    drp, dci, dv = (_t(dev, a) for a in (rp, ci, v))
    got = K.csr2csc(m, n, drp, dci, dv)
    torch.cuda.synchronize()
    for g, r in zip(got, ref):
        assert np.array_equal(g.cpu().numpy(), r)
    back = K.csc2csr(n, m, *got)
    ref2 = oracle.csc2csr(n, m, *ref)
    for g, r in zip(back, ref2):
        assert np.array_equal(g.cpu().numpy(), r)
    assert np.array_equal(back[0].cpu().numpy(), rp)


@pytest.mark.gpu
def test_csr2csc_config_sizes(sblas, dev):
    """At BASELINE size (config 2: 4M rows, 21M nnz): CSR -> CSC -> CSR returns the input (its rows are sorted),
    and the CSC of the symmetric Laplacian pattern equals its CSR."""
    import torch
    K = sblas.kernels
    rows, cols, rp, ci, v = sblas.synth.laplace2d(2048, 2048, torch.float64, dev)
    cp, ri, cv = K.csr2csc(rows, cols, rp, ci, v)
    assert torch.equal(cp, rp) and torch.equal(ri, ci) and torch.equal(cv, v)   # A = A^T, sorted rows
    rp2, ci2, v2 = K.csc2csr(cols, rows, cp, ri, cv)
    assert torch.equal(rp2, rp) and torch.equal(ci2, ci) and torch.equal(v2, v)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_coo_sort_and_loader(sblas, dev, oracle, golden, dtype):
    """CooSparseMatrix(const char*) (matrix.h:142-237): entries as the file holds them, then sorted by (row, col)."""
    import torch
    K = sblas.kernels
    h, w, r, c, v = K.read_mtx_coo(golden["ash85_path"])
    assert (h, w, len(r)) == (85, 85, 304)          # stored entries of the symmetric file, not expanded
    rng = np.random.default_rng(9)
    r = np.concatenate([r, rng.integers(0, h, 5000).astype(np.int32)])
    c = np.concatenate([c, rng.integers(0, w, 5000).astype(np.int32)])
    v = rng.standard_normal(len(r)).astype(dtype)
    rs, cs, vs = oracle.coo_sort(r, c, v.astype(np.float64))
    dr, dc, dv = (_t(dev, a) for a in (r, c, v))
    K.coo_sort(h, dr, dc, dv)
    torch.cuda.synchronize()
    assert np.array_equal(dr.cpu().numpy(), rs) and np.array_equal(dc.cpu().numpy(), cs)
    assert np.array_equal(dv.cpu().numpy(), vs.astype(dtype))


@pytest.mark.gpu
def test_spmv_auto_gather_heavy(sblas, dev, oracle):
    """AUTO on an input whose x does not fit L2: uniform-random columns switch to the sub-warp-per-row kernel
    (plan statistic col_jump > cols/256), banded columns stay on the staged kernel; both match the oracle."""
    import torch
    K = sblas.kernels
    rows, cols, deg = 20000, 14_000_000, 8
    rng = np.random.default_rng(12)
    rp = (np.arange(rows + 1, dtype=np.int64) * deg).astype(np.int32)
    x = rng.random(cols)
    y0 = rng.random(rows)
    for kind in ("random", "banded"):
        if kind == "random":
            ci = np.sort(rng.integers(0, cols, (rows, deg)), axis=1).astype(np.int32).reshape(-1)
        else:
            ci = (np.arange(rows)[:, None] * 3 + np.arange(deg)[None, :]).astype(np.int32).reshape(-1)
        v = rng.standard_normal(rows * deg)
        ref = oracle.spmv(rp, ci, v, x, y0.copy(), 1.5, -0.5)
        drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
        plan = K.Plan(drp, len(ci))
        assert plan.col_jump() == -1.0
        K.spmv(drp, dci, dv, dx, dy, 1.5, -0.5, plan=plan, cols=cols)
        torch.cuda.synchronize()
        cj = plan.col_jump()
        assert (cj > cols / 256.0) == (kind == "random"), (kind, cj)
        bound = _row_bound(oracle, rp, ci, v, x, 1.5, -0.5, y0)
        _check(dy.cpu().numpy(), ref, bound, np.float64)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("mat", ["short", "mid", "long", "empty_tail", "tiny"])
@pytest.mark.parametrize("panels", [1, 2, 3, 7, 32])
@pytest.mark.parametrize("fmt", ["sell", "csr"])
def test_spmv_column_panels(sblas, dev, oracle, dtype, mat, panels, fmt, monkeypatch):
    """SBLAS_SPMV_COLPANEL (colpanel.cu): one PIPE pass per column panel, the raw running row sums carried from
    pass to pass.  Within the row tolerance for every shape; with one lane per row (lanes = 1) and ascending
    columns the accumulation order is the CPU loop's, so the result is bit-identical to spmv.h:24-30."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(panels * 31 + len(mat))
    kw = MATS[mat]
    rp, ci, v = _rand_csr(rng, kw["h"], kw["w"], kw["avg"], kw["empty"], None, dtype)
    h, w = kw["h"], kw["w"]
    x, y0 = rng.random(w).astype(dtype), rng.random(h).astype(dtype)
    alpha, beta = dtype(1.25), dtype(-0.75)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), alpha, beta)
    drp, dci, dv, dx = (_t(dev, a) for a in (rp, ci, v, x))
    plan = K.Plan(drp, len(ci))
    # panels kept as sliced-ELL (32-row slices, rows sorted by length in windows of 256; taken when padding stays
    # below 25 %) or as CSR panels run by the staged kernel
    monkeypatch.setenv("SBLAS_COLPANEL_SELL", "1" if fmt == "sell" else "0")
    assert plan.build_colpanels(dci, dv, w) == (0, 0)          # automatic: x is tiny, no mirror
    with pytest.raises(sblas.SblasError, match="COLPANEL"):
        K.spmv(drp, dci, dv, dx, _t(dev, y0), alpha, beta, plan=plan, opts=sblas.make_opts(variant=sblas.SPMV_COLPANEL))
    n, W = plan.build_colpanels(dci, dv, w, n_panels=panels)
    assert n == panels and W % 32 == 0 and n * W >= w
    is_sell, stored = plan.colpanel_format()
    assert stored >= len(ci) and (fmt == "sell" or not is_sell)
    if is_sell:
        assert stored * 4 <= len(ci) * 5 + h
    bound = _row_bound(oracle, rp, ci, v, x, alpha, beta, y0)
    for lanes in (0, 1, 8):
        dy = _t(dev, y0)
        rep = _t(dev, np.full(h, -3.0, dtype))
        K.spmv(drp, dci, dv, dx, dy, alpha, beta, plan=plan,      # AUTO takes the mirror once the plan holds it
               opts=sblas.make_opts(lanes=lanes, replicas=[rep.data_ptr()]))
        torch.cuda.synchronize()
        got = dy.cpu().numpy()
        # one lane per row walks it in CSR order: bit-identical (sliced-ELL: for rows of <= 128 nnz per panel, longer ones are cut)
        longest = int((rp[1:] - rp[:-1]).max()) if h else 0
        _check(got, ref, bound, dtype, exact=(longest <= 128 if is_sell else lanes == 1))
        assert np.array_equal(rep.cpu().numpy(), got)
    # the plain kernel is still there on request, and agrees
    dy = _t(dev, y0)
    K.spmv(drp, dci, dv, dx, dy, alpha, beta, plan=plan, opts=sblas.make_opts(variant=sblas.SPMV_PIPE))
    _check(dy.cpu().numpy(), ref, bound, dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_spmv_column_panels_sliced_ell(sblas, dev, oracle, dtype, monkeypatch):
    """Even row lengths (the shape the mirror is built for): the panels are kept in sliced-ELL form, the result
    is bit-identical to the CPU loop (every lane walks its row in storage order), rows % 256 != 0."""
    import torch
    K = sblas.kernels
    monkeypatch.setenv("SBLAS_COLPANEL_SELL", "1")
    rng = np.random.default_rng(21)
    h, w, deg = 5000 + 77, 60000, 32
    rp = (np.arange(h + 1, dtype=np.int64) * deg).astype(np.int32)
    ci = np.sort(rng.integers(0, w, (h, deg)), axis=1).astype(np.int32).reshape(-1)
    v = (rng.random(h * deg) * 2 - 1).astype(dtype)
    x, y0 = rng.random(w).astype(dtype), rng.random(h).astype(dtype)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), dtype(0.5), dtype(2.0))
    drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
    plan = K.Plan(drp, len(ci))
    assert plan.build_colpanels(dci, dv, w, n_panels=4)[0] == 4
    is_sell, stored = plan.colpanel_format()
    assert is_sell and len(ci) <= stored <= 1.25 * len(ci) + h
    K.spmv(drp, dci, dv, dx, dy, 0.5, 2.0, plan=plan)
    torch.cuda.synchronize()
    assert np.array_equal(dy.cpu().numpy(), ref)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_spmv_virtual_row_sell_on_skewed_rows(sblas, dev, oracle, dtype, monkeypatch):
    """Power-law rows (R-MAT scale 17: rows of thousands of nnz beside empty ones) do not fit the plain sliced-ELL form;
    the automatic build keeps the VIRTUAL-ROW form instead (rows cut into pieces of 128, sorted by length in windows of
    65 536, a fix-up kernel adds the pieces): within the row tolerance, replicas written, HEAD / TAIL partial rows
    (row 0 is the biggest hub, i.e. a cut row) go to the carry buffer raw; SBLAS_SPMV_VSELL=0 keeps SPLIT."""
    import torch
    K, S = sblas.kernels, sblas.synth
    td = torch.float64 if dtype == np.float64 else torch.float32
    rows, cols, drp, dci, dv = S.rmat(17, 16, td, dev)
    rp, ci, v = drp.cpu().numpy(), dci.cpu().numpy(), dv.cpu().numpy()
    nnz = len(ci)
    assert nnz >= (1 << 20)
    rng = np.random.default_rng(3)
    x, y0 = rng.random(cols).astype(dtype), rng.random(rows).astype(dtype)
    dx = _t(dev, x)
    monkeypatch.setenv("SBLAS_SPMV_VSELL", "0")
    p0 = K.Plan(drp, nnz).prepare(dci, cols, v.itemsize)
    assert p0.build_colpanels(dci, dv, cols) == (0, 0)
    monkeypatch.setenv("SBLAS_SPMV_VSELL", "1")
    plan = K.Plan(drp, nnz).prepare(dci, cols, v.itemsize)
    assert plan.stats()["max_row_nnz"] > 1024
    n, W = plan.build_colpanels(dci, dv, cols)
    is_sell, stored = plan.colpanel_format()
    assert n == 1 and is_sell and nnz <= stored <= 1.05 * nnz + rows, (n, is_sell, stored, nnz)
    alpha, beta = dtype(1.25), dtype(-0.75)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), alpha, beta)
    raw = oracle.spmv(rp, ci, v, x, np.zeros(rows, dtype), dtype(1), dtype(0))
    bound = _row_bound(oracle, rp, ci, v, x, alpha, beta, y0)
    dy, rep = _t(dev, y0), _t(dev, np.full(rows, -3.0, dtype))
    # with a replica to write AUTO keeps SPLIT for a skewed input (sorted-order remote stores do not pay); the mirror
    # is still taken on request, and by AUTO when there is no replica
    K.spmv(drp, dci, dv, dx, dy, alpha, beta, plan=plan,
           opts=sblas.make_opts(variant=sblas.SPMV_COLPANEL, replicas=[rep.data_ptr()]))
    torch.cuda.synchronize()
    got = dy.cpu().numpy()
    _check(got, ref, bound, dtype)
    assert np.array_equal(rep.cpu().numpy(), got)
    short = (rp[1:] - rp[:-1]) <= 128          # uncut rows: one lane, storage order -> bit-identical
    assert np.array_equal(got[short], ref[short])
    dya = _t(dev, y0)
    K.spmv(drp, dci, dv, dx, dya, alpha, beta, plan=plan)                     # AUTO, no replica: the mirror
    assert np.array_equal(dya.cpu().numpy(), got)
    dyb, repb = _t(dev, y0), _t(dev, np.full(rows, -3.0, dtype))
    K.spmv(drp, dci, dv, dx, dyb, alpha, beta, plan=plan, opts=sblas.make_opts(replicas=[repb.data_ptr()]))   # AUTO + replica: SPLIT
    torch.cuda.synchronize()
    _check(dyb.cpu().numpy(), ref, bound, dtype)
    assert np.array_equal(repb.cpu().numpy(), dyb.cpu().numpy())
    # a second call on the same plan (the piece buffer is reused), with partial first / last rows
    dy2 = _t(dev, y0)
    carry = torch.full((8,), -1.0, dtype=td, device=dev)
    K.spmv(drp, dci, dv, dx, dy2, alpha, beta, plan=plan,
           opts=sblas.make_opts(flags=sblas.FLAG_HEAD_PARTIAL | sblas.FLAG_TAIL_PARTIAL, carry=carry.data_ptr()))
    torch.cuda.synchronize()
    y2, c = dy2.cpu().numpy(), carry.cpu().numpy()
    assert np.array_equal(y2[1:-1], got[1:-1]) and y2[0] == y0[0] and y2[-1] == y0[-1]
    rb = _row_bound(oracle, rp, ci, v, x, dtype(1), dtype(0), np.zeros(rows, dtype))
    assert abs(c[0] - raw[0]) <= TOL[np.dtype(dtype)] * rb[0] and abs(c[4] - raw[-1]) <= TOL[np.dtype(dtype)] * rb[-1]


@pytest.mark.gpu
def test_spmv_column_panels_partials(sblas, dev, oracle):
    """HEAD / TAIL partial rows of a multi-GPU segment go to the carry buffer from the LAST panel pass (raw sums)."""
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(4)
    h, w = 700, 5000
    rp, ci, v = _rand_csr(rng, h, w, 12, 0.0, None, np.float64)
    x, y0 = rng.random(w), rng.random(h)
    raw = oracle.spmv(rp, ci, v, x, np.zeros(h), 1.0, 0.0)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), 2.0, 3.0)
    drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
    carry = torch.full((8,), -1.0, dtype=torch.float64, device=dev)
    plan = K.Plan(drp, len(ci))
    plan.build_colpanels(dci, dv, w, n_panels=5)
    K.spmv(drp, dci, dv, dx, dy, 2.0, 3.0, plan=plan,
           opts=sblas.make_opts(lanes=1, flags=sblas.FLAG_HEAD_PARTIAL | sblas.FLAG_TAIL_PARTIAL, carry=carry.data_ptr()))
    torch.cuda.synchronize()
    y, c = dy.cpu().numpy(), carry.cpu().numpy()
    assert np.array_equal(y[1:-1], ref[1:-1]) and y[0] == y0[0] and y[-1] == y0[-1]
    assert c[0] == raw[0] and c[4] == raw[-1]


@pytest.mark.gpu
def test_spmv_column_panels_auto_on_gather_heavy(sblas, dev, oracle):
    """Automatic build (what sync2gpu asks for): an x above 96 MB with uniform-random columns gets the mirror
    (x slices of at most 48 MB), banded columns do not; AUTO then runs the panels and matches the oracle."""
    import torch
    K = sblas.kernels
    rows, cols, deg = 20000, 14_000_000, 8
    rng = np.random.default_rng(13)
    rp = (np.arange(rows + 1, dtype=np.int64) * deg).astype(np.int32)
    x, y0 = rng.random(cols), rng.random(rows)
    for kind in ("random", "banded"):
        if kind == "random":
            ci = np.sort(rng.integers(0, cols, (rows, deg)), axis=1).astype(np.int32).reshape(-1)
        else:
            ci = (np.arange(rows)[:, None] * 3 + np.arange(deg)[None, :]).astype(np.int32).reshape(-1)
        v = rng.standard_normal(rows * deg)
        ref = oracle.spmv(rp, ci, v, x, y0.copy(), 1.5, -0.5)
        drp, dci, dv, dx, dy = (_t(dev, a) for a in (rp, ci, v, x, y0))
        plan = K.Plan(drp, len(ci)).prepare(dci, cols, 8)
        assert (plan.col_jump() > cols / 256.0) == (kind == "random")
        n, W = plan.build_colpanels(dci, dv, cols)
        assert (n, W) == ((3, 4666688) if kind == "random" else (0, 0)), (kind, n, W)
        K.spmv(drp, dci, dv, dx, dy, 1.5, -0.5, plan=plan, cols=cols)
        torch.cuda.synchronize()
        _check(dy.cpu().numpy(), ref, _row_bound(oracle, rp, ci, v, x, 1.5, -0.5, y0), np.float64)


@pytest.mark.gpu
def test_python_container_mirrors(sblas, oracle, golden):
    """The host mirror's CooSparseMatrix / CscSparseMatrix / CsrToCsc / CscToCsr (unit_test.cu:20-51)."""
    A = sblas.CsrSparseMatrix(golden["ash85_path"])
    csc = sblas.CscSparseMatrix(A)
    cp, ri, cv = oracle.csr2csc(A.height, A.width, A.csrRowPtr, A.csrColIdx, A.csrVal)
    assert np.array_equal(csc.cscColPtr, cp) and np.array_equal(csc.cscRowIdx, ri) and np.array_equal(csc.cscVal, cv)
    ci2, rp2, v2 = sblas.CscToCsr(A.width, A.height, A.nnz, csc.cscColPtr, csc.cscRowIdx, csc.cscVal)
    assert np.array_equal(rp2, A.csrRowPtr) and all(
        np.all(np.diff(ci2[rp2[r]:rp2[r + 1]]) >= 0) for r in range(A.height))
    coo = sblas.CooSparseMatrix(golden["ash85_path"], 2, sblas.segment)
    assert coo.nnz == 304 and list(coo.nnz_gpu) == [152, 152]
    key = coo.cooRowIdx.astype(np.int64) * coo.width + coo.cooColIdx
    assert np.all(np.diff(key) >= 0)
    assert np.array_equal(coo.cooRowIdx_gpu[1].cpu().numpy(), coo.cooRowIdx[152:])
