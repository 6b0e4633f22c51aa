"""Parity at the sizes BASELINE.json quotes (missing #4): the AUTO kernels on the full-size
configs 2 and 4, on configs 3 and 5 scaled down by 8, and on the two 16M-row x 32 SpMV shapes (uniform-random:
column-panel mirror; banded: the consecutive-tile order of the staged kernel), each checked against the oracle
on row windows spread over the matrix (first rows, 71 %, last rows: nnz offsets beyond 2^30, large grids) and,
on ALL rows, against a size-independent property (an independent kernel for SpMV, linearity for SpMM)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = {"f64": 1e-12, "f32": 1e-5}


def _windows(rows, n):
    out = []
    for f in (0.0, 0.71, 1.0):
        a = max(0, min(int(rows * f), rows - n))
        if all(a >= b + n or a + n <= b for b in out):
            out.append(a)
    return out


def _csr_window(rp, ci, v, a, n):
    rps = rp[a:a + n + 1].cpu().numpy().astype(np.int64)
    n0, n1 = int(rps[0]), int(rps[-1])
    return (rps - n0).astype(np.int32), ci[n0:n1].cpu().numpy(), v[n0:n1].cpu().numpy()


def _spmv_case(sblas, dev, oracle, gen, dtype, n_win, expect=None):
    import torch
    K, S = sblas.kernels, sblas.synth
    td = torch.float64 if dtype == "f64" else torch.float32
    rows, cols, rp, ci, v = gen(td)
    x = S.fill_uniform(torch.empty(cols, dtype=td, device=dev), stream_id=11)
    y = torch.ones(rows, dtype=td, device=dev)
    plan = K.Plan(rp, v.numel()).prepare(ci, cols, v.element_size())
    panels = plan.build_colpanels(ci, v, cols)          # automatic, as sync2gpu asks for it
    K.spmv(rp, ci, v, x, y, 1.0, 1.0, plan=plan, cols=cols)
    torch.cuda.synchronize()
    if expect is not None:
        expect(plan, panels)
    xh = x.cpu().numpy()
    npdt = np.float64 if dtype == "f64" else np.float32
    for a in _windows(rows, n_win):
        rps, cis, vs = _csr_window(rp, ci, v, a, min(n_win, rows - a))
        R = len(rps) - 1
        ref = oracle.spmv(rps, cis, vs, xh, np.ones(R, npdt), npdt(1), npdt(1))
        bound = oracle.absrow_spmv(rps, cis, vs, xh) + 1.0
        err = np.abs(y[a:a + R].cpu().numpy().astype(np.float64) - ref.astype(np.float64)) / bound
        assert err.max() <= TOL[dtype], "rows [%d,%d): %g" % (a, a + R, err.max())
    # all rows: the independent sub-warp-per-row kernel
    y2 = torch.ones(rows, dtype=td, device=dev)
    K.spmv(rp, ci, v, x, y2, 1.0, 1.0, opts=sblas.make_opts(variant=sblas.SPMV_VECTOR), cols=cols)
    bnd = torch.ones(rows, dtype=td, device=dev)
    K.spmv(rp, ci, v.abs(), x, bnd, 1.0, 1.0, opts=sblas.make_opts(variant=sblas.SPMV_VECTOR), cols=cols)
    assert float(((y - y2).abs() / bnd).max()) <= 2 * TOL[dtype]


@pytest.mark.parametrize("mirror", ["1", "0"])
def test_c2_full_size(sblas, dev, oracle, mirror, monkeypatch):
    """configs[1]: 2-D Laplacian, 4M rows, fp64 SpMV, AUTO: the single sliced-ELL panel sync2gpu keeps for regular
    rows (4 % padding: boundary rows have 3-4 nnz), and without mirrors the staged PIPE kernel, strided tile order."""
    monkeypatch.setenv("SBLAS_SPMV_MIRROR", mirror)

    def expect(plan, panels):
        assert panels == ((1, 1 << 22) if mirror == "1" else (0, 0)), panels
    _spmv_case(sblas, dev, oracle, lambda td: sblas.synth.laplace2d(2048, 2048, td, dev), "f64", 20000, expect)


@pytest.mark.parametrize("vsell", ["1", "0"])
def test_c4_full_size(sblas, dev, oracle, vsell, monkeypatch):
    """configs[3]: R-MAT scale 23, fp32 SpMV, AUTO: the virtual-row sliced-ELL mirror (rows of 10^5 nnz cut into pieces
    of 128) and, without it, SPLIT (rows cut across CTAs); the first window holds the hub rows."""
    monkeypatch.setenv("SBLAS_SPMV_VSELL", vsell)

    def expect(plan, panels):
        assert panels[0] == (1 if vsell == "1" else 0), panels
    _spmv_case(sblas, dev, oracle, lambda td: sblas.synth.rmat(23, 16, td, dev), "f32", 6000, expect)


def test_c3v_full_size_column_panels(sblas, dev, oracle):
    """the north star's SpMV shape: uniform-random 16M x 32, fp64; x (134 MB) does not fit L2, sync2gpu builds the
    column-panel mirror (3 panels of 43 MB of x, sliced-ELL) and AUTO runs it; nnz offsets reach 2^29."""
    def expect(plan, panels):
        assert plan.col_jump() > (1 << 24) / 256.0
        assert panels == (3, 5592416), panels
        is_sell, stored = plan.colpanel_format()
        assert is_sell and stored <= 1.25 * (1 << 29)
    _spmv_case(sblas, dev, oracle, lambda td: sblas.synth.uniform(1 << 24, 1 << 24, 32, td, dev), "f64", 4096, expect)


def test_b32_full_size_single_panel_sell(sblas, dev, oracle):
    """banded 16M x 32, fp64: local rows (col_jump ~ 1), regular row lengths -> sync2gpu keeps ONE sliced-ELL panel and
    AUTO runs the per-lane kernel on it (no padding: every row has 32 nnz)."""
    def expect(plan, panels):
        assert 0.0 <= plan.col_jump() < 4.0 and panels == (1, 1 << 24)
        is_sell, stored = plan.colpanel_format()
        assert is_sell and stored == (1 << 29)
    _spmv_case(sblas, dev, oracle, lambda td: sblas.synth.banded(1 << 24, 32, td, dev), "f64", 4096, expect)


def test_b32_full_size_consecutive_tiles(sblas, dev, oracle, monkeypatch):
    """the same input without any mirror (SBLAS_SPMV_MIRROR=0): x larger than L2 but local rows -> the staged PIPE
    kernel with CONSECUTIVE tiles per CTA (spmv.cu launch_pipe: rows * col_jump * 8 B > 96 MB)."""
    monkeypatch.setenv("SBLAS_SPMV_MIRROR", "0")

    def expect(plan, panels):
        assert 0.0 <= plan.col_jump() < 4.0 and panels == (0, 0)
    _spmv_case(sblas, dev, oracle, lambda td: sblas.synth.banded(1 << 24, 32, td, dev), "f64", 4096, expect)


def _spmm_case(sblas, dev, oracle, gen, k, n_win, flags=0):
    import torch
    K, S = sblas.kernels, sblas.synth
    td = torch.float64
    rows, cols, rp, ci, v = gen(td)
    B = S.fill_uniform(torch.empty(cols * k, dtype=td, device=dev), stream_id=11)     # row-major [cols x k]
    C = torch.ones(rows * k, dtype=td, device=dev)
    plan = K.Plan(rp, v.numel())
    plan.build_rowblocks(ci, v)
    K.spmm(rp, ci, v, k, sblas.ROW_MAJOR, B, k, C, k, 1.0, 1.0, cols, plan=plan, opts=sblas.make_opts(flags=flags))
    torch.cuda.synchronize()
    Bc = np.ascontiguousarray(B.view(cols, k).t().cpu().numpy()).reshape(-1)           # col-major host operand
    for a in _windows(rows, n_win):
        rps, cis, vs = _csr_window(rp, ci, v, a, min(n_win, rows - a))
        R = len(rps) - 1
        ref = oracle.spmm(rps, cis, vs, k, Bc, np.ones(R * k), 1.0, 1.0).reshape(k, R)
        got = C.view(rows, k)[a:a + R].cpu().numpy()
        if flags & sblas.FLAG_EXACT:
            assert np.array_equal(np.ascontiguousarray(got.T), ref)
        for n in range(0, k, max(1, k // 16)):
            bound = oracle.absrow_spmv(rps, cis, vs, Bc[n * cols:(n + 1) * cols]) + 1.0
            assert (np.abs(got[:, n] - ref[n]) / bound).max() <= 1e-12
    # all rows: (C0 + A B).1 = C0.1 + A (B.1)
    xs = B.view(cols, k).sum(1).contiguous()
    ys = torch.full((rows,), float(k), dtype=td, device=dev)
    K.spmv(rp, ci, v, xs, ys, 1.0, 1.0, cols=cols)
    bnd = torch.full((rows,), float(k), dtype=td, device=dev)
    K.spmv(rp, ci, v.abs(), xs, bnd, 1.0, 1.0, cols=cols)
    assert float(((C.view(rows, k).sum(1) - ys).abs() / bnd).max()) <= 2e-12


@pytest.mark.parametrize("flags", ["default", "exact"])
def test_c3_scaled_by_8(sblas, dev, oracle, flags):
    """configs[2] / 8: uniform-random 2M x 32, fp64 SpMM k = 64 (CSR ROWMAJOR kernel; the mirror's fill is 0.26)."""
    fl = sblas.FLAG_EXACT if flags == "exact" else 0
    _spmm_case(sblas, dev, oracle, lambda td: sblas.synth.uniform(1 << 21, 1 << 21, 32, td, dev), 64, 768, fl)


@pytest.mark.parametrize("flags", ["default", "exact"])
def test_c5_scaled_by_8(sblas, dev, oracle, flags):
    """configs[4] / 8: banded 256K x 64, fp64 SpMM k = 256 (row-block pipe kernel, two 128-column panels)."""
    fl = sblas.FLAG_EXACT if flags == "exact" else 0
    _spmm_case(sblas, dev, oracle, lambda td: sblas.synth.banded(1 << 18, 64, td, dev), 256, 256, fl)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1), (1000, 3), (70001, 64), (5, 700), (300000, 17)])
def test_streamed_dense_rows_roundtrip(sblas, dev, dtype, shape):
    """sblas_dense_upload_rows_* / _download_rows_* (the streamed sync2gpu / sync2cpu of a row range): col-major
    pinned host matrix <-> row-major device block, chunks of rows, copy overlapping conversion."""
    import torch
    K = sblas.kernels
    rows, k = shape
    rng = np.random.default_rng(rows + k)
    td = torch.float64 if dtype == np.float64 else torch.float32
    h = torch.from_numpy(rng.random(rows * k).astype(dtype)).pin_memory()          # col-major [k][rows]
    ld = k + 3
    r0, nr = rows // 3, rows - rows // 3 - rows // 5
    d = torch.full((max(nr, 1) * ld,), -1.0, dtype=td, device=dev)
    K.dense_upload_rows(h, rows, r0, nr, k, d, ld)
    torch.cuda.synchronize()
    want = h.view(k, rows)[:, r0:r0 + nr].t()
    got = d.view(-1, ld)[:nr]
    assert torch.equal(got[:, :k].cpu(), want.contiguous()) and bool((got[:, k:] == -1.0).all())
    back = torch.full((rows * k,), -2.0, dtype=td).pin_memory()
    K.dense_download_rows(d, ld, nr, k, back, rows, r0)
    torch.cuda.synchronize()
    bv = back.view(k, rows)
    assert torch.equal(bv[:, r0:r0 + nr], h.view(k, rows)[:, r0:r0 + nr])
    assert bool((bv[:, :r0] == -2.0).all()) and bool((bv[:, r0 + nr:] == -2.0).all())
