"""The reference's own end-to-end tests, re-created against the drop-in operator interface:
spmvCsrTest / spmmCsrTest / spmmCsrTest2 (unit_test.cu:66-175, spmv_test.cu:6-42,
spmm_test.cu:19-93): build A from ash85.mtx, distribute with the policies the reference
uses, run the multi-GPU op, compare with the CPU function.  With fewer physical devices than
n_gpu the logical GPUs are folded onto the available ones (SBLAS_DEVICE_MAP, conftest.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact_spmm(sblas):
    """These tests prove bit-identity with the reference's CPU loops: SpMM runs with separately rounded
    mul/add (sblas_set_spmm_exact).  The default fused multiply-add arithmetic is covered by the
    tolerance tests, which opt out with SBLAS_FLAG_FMA or switch the default back."""
    prev = sblas.kernels.set_spmm_exact(True)
    yield
    sblas.kernels.set_spmm_exact(prev)


def _bound(oracle, A, x, alpha, beta, y0):
    return abs(alpha) * oracle.absrow_spmv(A.csrRowPtr, A.csrColIdx, A.csrVal, x) + np.abs(beta * y0) + 1e-300


@pytest.mark.parametrize("n_gpu", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("alpha,beta", [(1.0, 1.0), (3.0, 4.0)])
def test_spmv_v1_ash85(sblas, dev, oracle, golden, n_gpu, alpha, beta):
    """spmv_test.cu:6-42 / unit_test.cu:185 (alpha=3, beta=4, 4 GPUs)."""
    S = sblas
    A = S.CsrSparseMatrix(golden["ash85_path"])
    B = S.DenseVector(A.width, 1.0)
    C = S.DenseVector(A.height, 1.0)
    A.sync2gpu(n_gpu, S.segment, quiet=True)
    B.sync2gpu(n_gpu, S.replicate)
    C.sync2gpu(n_gpu, S.replicate)
    g = golden["meta"]["ash85"]["partition"][str(n_gpu)]
    assert A.nnz_gpu.tolist() == g["nnz_gpu"] and A.starting_row_gpu.tolist() == g["start_row"]
    for i in range(n_gpu):   # bit-exact local row pointers on the device (matrix.h:372-375)
        assert np.array_equal(A.csrRowPtr_gpu[i].cpu().numpy(), golden["ash_part%d_rowptr%d" % (n_gpu, i)])
    S.sblas_spmv_csr_v1(A, B, C, alpha, beta, n_gpu)
    ref = oracle.spmv(A.csrRowPtr, A.csrColIdx, A.csrVal, np.ones(85), np.ones(85), alpha, beta)
    for i in range(n_gpu):   # every replica is updated (matrix.h:714-726)
        C.sync2cpu(i)
        assert np.array_equal(C.val, ref)   # ash85 is all-ones: sums are exact integers
        assert oracle.check_equal(ref, C.val)


@pytest.mark.parametrize("n_gpu", [1, 2, 4, 8])
@pytest.mark.parametrize("k,alpha,beta", [(64, 1.0, 1.0), (256, 3.0, 4.0)])
@pytest.mark.parametrize("layout", ["row", "col"])
def test_spmm_v1_ash85(sblas, dev, oracle, golden, n_gpu, k, alpha, beta, layout):
    """Method 1, spmm_test.cu:19-56 / README.md:76 (`./spmm 1 ./ash85.mtx 64 1.0 1.0 4`)."""
    S = sblas
    A = S.CsrSparseMatrix(golden["ash85_path"])
    B = S.DenseMatrix(A.width, k, data=oracle.rand_fill(A.width * k))     # seed-211 ctor
    C = S.DenseMatrix(A.height, k, 1.0)
    A.sync2gpu(n_gpu, S.replicate)
    B.sync2gpu(n_gpu, S.segment, device_layout=layout)
    C.sync2gpu(n_gpu, S.segment, device_layout=layout)
    assert B.dim_gpu.tolist() == golden["meta"]["dense_split"]["%d/%d" % (k, n_gpu)][1]
    S.sblas_spmm_csr_v1(A, B, C, alpha, beta, n_gpu)     # refreshes the host copy (spmm.h:151)
    ref = golden["ash_spmm_k%d" % k]
    if layout == "row":
        assert np.array_equal(C.val, ref)                # sequential per (row, col): bit-identical
    else:
        assert np.abs(C.val - ref).max() <= 1e-12 * np.abs(ref).max()
    assert oracle.check_equal(ref, C.val)


@pytest.mark.parametrize("n_gpu", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("k,alpha,beta", [(64, 1.0, 1.0), (256, 3.0, 4.0), (9, 0.5, 0.0)])
@pytest.mark.parametrize("layout", ["row", "col"])
def test_spmm_v2_ash85(sblas, dev, oracle, golden, n_gpu, k, alpha, beta, layout):
    """Method 2, spmm_test.cu:58-93 / unit_test.cu:184."""
    S = sblas
    A = S.CsrSparseMatrix(golden["ash85_path"])
    B = S.DenseMatrix(A.width, k, data=oracle.rand_fill(A.width * k))
    C = S.DenseMatrix(A.height, k, 1.0)
    A.sync2gpu(n_gpu, S.segment, quiet=True)
    B.sync2gpu(n_gpu, S.replicate, device_layout=layout)
    C.sync2gpu(n_gpu, S.replicate, device_layout=layout)
    S.sblas_spmm_csr_v2(A, B, C, alpha, beta, n_gpu)
    ref = golden["ash_spmm_k%d" % k]
    for i in range(n_gpu):
        C.sync2cpu(i)
        assert np.abs(C.val - ref).max() <= 1e-12 * np.abs(ref).max(), i
        assert oracle.check_equal(ref, C.val)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n_gpu", [2, 4, 7])
def test_rowsplit_hard_partitions(sblas, dev, oracle, dtype, n_gpu):
    """A row spanning several GPUs, empty rows before the first / after the last segment and
    between segments: the cases the reference handles through its zero buffer + AllReduce."""
    S = sblas
    rng = np.random.default_rng(n_gpu)
    h = w = 400
    lens = rng.poisson(3, h)
    lens[:17] = 0            # leading empty rows (never covered by a segment)
    lens[-29:] = 0           # trailing empty rows
    lens[100:140] = 0
    lens[200] = 390          # spans several segments
    rp = np.zeros(h + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    ci = np.concatenate([np.sort(rng.choice(w, l, replace=False)) for l in lens]).astype(np.int32)
    v = (rng.random(rp[-1]) * 2 - 1).astype(dtype)
    x, y0 = rng.random(w).astype(dtype), rng.random(h).astype(dtype)
    tol = 1e-12 if dtype == np.float64 else 1e-5
    A = S.CsrSparseMatrix(height=h, width=w, rowptr=rp, colidx=ci, val=v)
    A.sync2gpu(n_gpu, S.segment, quiet=True)
    q = oracle.partition_nnz(rp, n_gpu, float_ceil=False)
    assert np.array_equal(A.starting_row_gpu, q["start_row"]) and np.array_equal(A.stoping_row_gpu, q["stop_row"])
    # SpMV
    B, C = S.DenseVector(w, data=x, dtype=dtype), S.DenseVector(h, data=y0, dtype=dtype)
    B.sync2gpu(n_gpu, S.replicate)
    C.sync2gpu(n_gpu, S.replicate)
    S.sblas_spmv_csr_v1(A, B, C, 1.5, -0.5, n_gpu)
    ref = oracle.spmv(rp, ci, v, x, y0.copy(), dtype(1.5), dtype(-0.5))
    bound = _bound(oracle, A, x, 1.5, -0.5, y0.astype(np.float64))
    for i in range(n_gpu):
        C.sync2cpu(i)
        assert (np.abs(C.val.astype(np.float64) - ref) / bound).max() <= tol
    # SpMM method 2, both device layouts
    k = 10
    Bc, C0 = rng.random(w * k).astype(dtype), rng.random(h * k).astype(dtype)
    refC = oracle.spmm(rp, ci, v, k, Bc, C0.copy(), dtype(1.5), dtype(-0.5))
    for layout in ("row", "col"):
        Bm, Cm = S.DenseMatrix(w, k, data=Bc, dtype=dtype), S.DenseMatrix(h, k, data=C0, dtype=dtype)
        Bm.sync2gpu(n_gpu, S.replicate, device_layout=layout)
        Cm.sync2gpu(n_gpu, S.replicate, device_layout=layout)
        S.sblas_spmm_csr_v2(A, Bm, Cm, 1.5, -0.5, n_gpu)
        for i in range(n_gpu):
            Cm.sync2cpu(i)
            for n in range(k):
                b = _bound(oracle, A, Bc[n * w:(n + 1) * w], 1.5, -0.5, C0[n * h:(n + 1) * h].astype(np.float64))
                err = np.abs(Cm.val[n * h:(n + 1) * h].astype(np.float64) - refC[n * h:(n + 1) * h]) / b
                assert err.max() <= tol, (layout, i, n)


def test_operator_preconditions(sblas, dev, golden):
    """Same error behaviour as the reference: assert on shape mismatch (spmv.h:40-41),
    exit(-1) on row-major operands (spmm.h:91-98)."""
    S = sblas
    A = S.CsrSparseMatrix(golden["ash85_path"])
    A.sync2gpu(1, S.segment, quiet=True)
    B, C = S.DenseVector(84, 1.0), S.DenseVector(85, 1.0)
    B.sync2gpu(1, S.replicate)
    C.sync2gpu(1, S.replicate)
    with pytest.raises(AssertionError):
        S.sblas_spmv_csr_v1(A, B, C, 1.0, 1.0, 1)
    Bm = S.DenseMatrix(85, 4, 1.0, order=S.row_major)
    Cm = S.DenseMatrix(85, 4, 1.0)
    with pytest.raises(SystemExit):
        S.sblas_spmm_csr_v2(A, Bm, Cm, 1.0, 1.0, 1)
    with pytest.raises(AssertionError):
        S.DenseVector(5, 1.0).sync2gpu(2, S.segment)      # matrix.h:693
