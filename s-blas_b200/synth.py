"""Synthetic CSR matrices / dense operands of BASELINE.json's configs, generated ON THE DEVICE
through the C ABI (sblas_synth_*).  Replaces matrices/get_matrices.sh (network) of the
reference.  tests/synth_ref.py restates every generator in numpy; tests compare bit for bit.
"""
import ctypes as C

from ._lib import check, lib, stream_ptr, vp

SEED = 211  # config.h:23 RAND_INIT_SEED


def _torch():
    import torch
    return torch


def _alloc_csr(rows, nnz, dtype, device):
    torch = _torch()
    # col/val padded to a multiple of 4 entries so 128-bit loads of the last group stay in bounds
    pad = (nnz + 3) // 4 * 4
    rp = torch.empty(rows + 1, dtype=torch.int32, device=device)
    ci = torch.zeros(pad, dtype=torch.int32, device=device)[:nnz]
    v = torch.zeros(pad, dtype=dtype, device=device)[:nnz]
    return rp, ci, v


def fill_uniform(t, seed=SEED, stream_id=0, lo=0.0, hi=1.0):
    """t[i] = lo + (hi-lo)*u01(seed, stream_id, i), in place."""
    torch = _torch()
    f64 = t.dtype == torch.float64
    fn = lib().sblas_synth_fill_f64 if f64 else lib().sblas_synth_fill_f32
    ct = C.c_double if f64 else C.c_float
    with torch.cuda.device(t.device):
        check(fn(stream_ptr(), vp(t), C.c_int64(t.numel()), C.c_uint64(seed), C.c_uint64(stream_id),
                 ct(lo), ct(hi)), "synth_fill")
    return t


def laplace2d(nx, ny, dtype, device):
    """2-D 5-point Laplacian (BASELINE config 2: nx = ny = 2048)."""
    torch = _torch()
    rows = nx * ny
    nnz = int(lib().sblas_synth_laplace2d_nnz(C.c_int32(nx), C.c_int32(ny)))
    rp, ci, v = _alloc_csr(rows, nnz, dtype, device)
    with torch.cuda.device(device):
        check(lib().sblas_synth_laplace2d(stream_ptr(), C.c_int32(nx), C.c_int32(ny), vp(rp), vp(ci),
                                          vp(v), C.c_int(dtype == torch.float64)), "laplace2d")
    return rows, rows, rp, ci, v


def uniform(rows, cols, deg, dtype, device, seed=SEED):
    """Exactly `deg` distinct uniform columns per row (BASELINE config 3: 2^24 x 2^24, 32)."""
    torch = _torch()
    rp, ci, v = _alloc_csr(rows, rows * deg, dtype, device)
    with torch.cuda.device(device):
        check(lib().sblas_synth_uniform(stream_ptr(), C.c_int32(rows), C.c_int32(cols),
                                        C.c_int32(deg), C.c_uint64(seed), vp(rp), vp(ci), vp(v),
                                        C.c_int(dtype == torch.float64)), "synth_uniform")
    return rows, cols, rp, ci, v


def banded(n, bw, dtype, device, seed=SEED):
    """Row i holds bw consecutive columns from clamp(i-bw/2, 0, n-bw) (BASELINE config 5)."""
    torch = _torch()
    rp, ci, v = _alloc_csr(n, n * bw, dtype, device)
    with torch.cuda.device(device):
        check(lib().sblas_synth_banded(stream_ptr(), C.c_int32(n), C.c_int32(bw), C.c_uint64(seed),
                                       vp(rp), vp(ci), vp(v), C.c_int(dtype == torch.float64)),
              "synth_banded")
    return n, n, rp, ci, v


def rmat(scale, edge_factor, dtype, device, seed=SEED):
    """R-MAT a,b,c,d = .57,.19,.19,.05, duplicates merged (BASELINE config 4: scale 23, ef 16)."""
    torch = _torch()
    rows = 1 << scale
    nnz = C.c_int64()
    with torch.cuda.device(device):
        check(lib().sblas_synth_rmat(stream_ptr(), C.c_int32(scale), C.c_int32(edge_factor),
                                     C.c_uint64(seed), C.byref(nnz), None, None, None, C.c_int(0)),
              "synth_rmat(count)")
        rp, ci, v = _alloc_csr(rows, nnz.value, dtype, device)
        check(lib().sblas_synth_rmat(stream_ptr(), C.c_int32(scale), C.c_int32(edge_factor),
                                     C.c_uint64(seed), C.byref(nnz), vp(rp), vp(ci), vp(v),
                                     C.c_int(dtype == torch.float64)), "synth_rmat(fill)")
    return rows, rows, rp, ci, v
