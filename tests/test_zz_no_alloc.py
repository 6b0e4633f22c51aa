"""sblas_spmv_csr_noalloc_f64 — the allocation-free SpMV entry point added
(no counterpart in the reference): its scratch is a fixed device array inside the library.  Collected last on
purpose.  Passes alone in a fresh process and after the other tests: it builds all
of its own inputs and depends on no state of the library."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12   # the fp64 row tolerance of the other parity tests, relative to |alpha| (|A||x|)_i + |beta y_i|


def _skewed_csr(rng, h, w, avg, long_rows):
    lens = rng.poisson(avg, h).astype(np.int64)
    lens[rng.random(h) < 0.2] = 0
    for r, n in long_rows:
        lens[r] = n
    lens = np.minimum(lens, w)
    rp = np.zeros(h + 1, np.int32)
    rp[1:] = np.cumsum(lens)
    ci = np.concatenate([np.sort(rng.choice(w, int(n), replace=False)) for n in lens] + [np.zeros(0, np.int64)]).astype(np.int32)
    v = rng.random(int(rp[-1])) * 2 - 1
    return rp, ci, v


def _free_bytes(torch):
    return torch.cuda.mem_get_info()[0]


@pytest.mark.parametrize("case", ["skewed", "short"])
def test_noalloc_spmv_matches_the_existing_entry_point(sblas, dev, oracle, case):
    import torch
    K = sblas.kernels
    rng = np.random.default_rng(7 if case == "skewed" else 8)
    if case == "skewed":   # rows cut across several tiles of 1536 merge items, empty rows, a long last row
        h, w = 6000, 40000
        rp, ci, v = _skewed_csr(rng, h, w, 6, [(17, 9000), (18, 5000), (h - 1, 3000)])
    else:
        h, w = 20000, 20000
        rp, ci, v = _skewed_csr(rng, h, w, 4, [])
    x, y0 = rng.random(w), rng.random(h)
    alpha, beta = 1.25, -0.75
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    drp, dci, dv, dx = t(rp), t(ci), t(v), t(x)
    # the existing entry point (sblas_spmv_csr_f64 with a plan) and the CPU oracle
    y_ref = t(y0)
    K.spmv(drp, dci, dv, dx, y_ref, alpha, beta, plan=K.Plan(drp, len(ci)))
    torch.cuda.synchronize()
    ref = y_ref.cpu().numpy()
    cpu = oracle.spmv(rp, ci, v, x, y0.copy(), alpha, beta)
    bound = abs(alpha) * oracle.absrow_spmv(rp, ci, v, x) + np.abs(beta * y0) + 1e-300
    # two calls in a row (the second sees the scratch the first left behind), on fresh outputs allocated up front
    outs = [t(y0), t(y0)]
    torch.cuda.synchronize()
    free0 = _free_bytes(torch)
    for y in outs:
        K.spmv_noalloc(drp, dci, dv, dx, y, alpha, beta)
    torch.cuda.synchronize()
    assert _free_bytes(torch) == free0, "the no-alloc entry point changed the device's free memory"
    for i, y in enumerate(outs):
        got = y.cpu().numpy()
        assert (np.abs(got - ref) / bound).max() <= TOL, "call %d vs sblas_spmv_csr_f64" % i
        assert (np.abs(got - cpu) / bound).max() <= TOL, "call %d vs the CPU oracle" % i
    assert np.array_equal(outs[0].cpu().numpy(), outs[1].cpu().numpy())
    # a different matrix right after (smaller: stale table entries of the larger call stay behind it)
    rp2, ci2, v2 = _skewed_csr(rng, 300, w, 3, [(5, 2000)])
    y2 = rng.random(300)
    d2 = [t(a) for a in (rp2, ci2, v2)]
    g2 = t(y2)
    K.spmv_noalloc(d2[0], d2[1], d2[2], dx, g2, alpha, beta)
    cpu2 = oracle.spmv(rp2, ci2, v2, x, y2.copy(), alpha, beta)
    b2 = abs(alpha) * oracle.absrow_spmv(rp2, ci2, v2, x) + np.abs(beta * y2) + 1e-300
    assert (np.abs(g2.cpu().numpy() - cpu2) / b2).max() <= TOL


def test_noalloc_spmv_rejects_sizes_beyond_its_scratch(sblas, dev):
    """rows + nnz beyond 131 072 tiles x 1536: a clear error before anything is launched (the pointers are never
    dereferenced, so small buffers do)."""
    import torch
    lib = sblas._lib.lib()
    z = torch.zeros(8, dtype=torch.float64, device=dev)
    zi = torch.zeros(8, dtype=torch.int32, device=dev)
    vp = lambda a: C.c_void_p(a.data_ptr())
    n0 = lib.sblas_launch_count()
    rc = lib.sblas_spmv_csr_noalloc_f64(None, C.c_int32(1 << 20), C.c_int32(8), C.c_int64(201_326_592), vp(zi), vp(zi), vp(z),
                                        vp(z), vp(z), C.c_double(1.0), C.c_double(0.0))
    assert rc != 0 and b"exceeds the fixed scratch" in lib.sblas_last_error()
    assert lib.sblas_launch_count() == n0
    with pytest.raises(sblas.SblasError, match="fixed scratch"):
        big = torch.zeros(2, dtype=torch.int32, device=dev)
        sblas._lib.check(lib.sblas_spmv_csr_noalloc_f64(None, C.c_int32(1), C.c_int32(8), C.c_int64((1 << 31) - 1), vp(big), vp(zi),
                                                        vp(z), vp(z), vp(z), C.c_double(1.0), C.c_double(0.0)), "spmv_noalloc")
