"""The reference's operator interface for the hot path, same names and argument meaning:

    sblas_spmv_csr_v1(pA, pB, pC, alpha, beta, n_gpu)      spmv.h:36-139
    sblas_spmm_csr_v1(pA, pB, pC, alpha, beta, n_gpu)      spmm.h:84-161   (method 1)
    sblas_spmm_csr_v2(pA, pB, pC, alpha, beta, n_gpu)      spmm.h:164-284  (method 2)

Pre-conditions are the reference's (operands already distributed with sync2gpu using the
policies its tests use, spmm_test.cu:28-30/:68-70, spmv_test.cu:16-18); violations raise
AssertionError where the reference asserts and SystemExit where it exit(-1)s.  Each op is one
call into the C ABI (sblas_mgpu_*); there is no CPU path.
"""
import ctypes as C
import sys

import numpy as np

from . import matrix as M
from ._lib import check, lib, COL_MAJOR, ROW_MAJOR


def _sfx(dt):
    return "f64" if np.dtype(dt) == np.float64 else "f32"


def _ptr_array(tensors):
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr() if t is not None and t.numel() > 0 else None
    return arr


def _plan_array(plans):
    arr = (C.c_void_p * len(plans))()
    for i, p in enumerate(plans):
        arr[i] = p.handle if p is not None else None
    return arr


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


def _sc(dt, v):
    return C.c_double(float(v)) if np.dtype(dt) == np.float64 else C.c_float(float(v))


def sblas_spmv_csr_v1(pA, pB, pC, alpha, beta, n_gpu):
    assert pA.width == pB.length                        # spmv.h:40
    assert pA.height == pC.length                       # spmv.h:41
    assert pA.policy == M.segment and pA.n_gpu == n_gpu, "A must be sync2gpu(n_gpu, segment)"
    assert pB.policy == M.replicate and pC.policy == M.replicate
    dt = pA.dtype
    ng, s, e = _i32(pA.nnz_gpu), _i32(pA.starting_row_gpu), _i32(pA.stoping_row_gpu)
    fn = getattr(lib(), "sblas_mgpu_spmv_csr_v1_" + _sfx(dt))
    check(fn(C.c_uint(n_gpu), C.c_int32(pA.height), C.c_int32(pA.width),
             _ptr_array(pA.csrRowPtr_gpu), _ptr_array(pA.csrColIdx_gpu), _ptr_array(pA.csrVal_gpu),
             ng.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p),
             e.ctypes.data_as(C.c_void_p), _ptr_array(pB.val_gpu), _ptr_array(pC.val_gpu),
             _sc(dt, alpha), _sc(dt, beta), _plan_array(pA.plans)), "sblas_spmv_csr_v1")


def _check_colmajor(pB, pC, name):
    if pB.order == M.row_major:                          # spmm.h:91-94
        sys.stderr.write("%s: B should be in column major!\n" % name)
        raise SystemExit(-1)
    if pC.order == M.row_major:                          # spmm.h:95-98
        sys.stderr.write("%s: C should be in column major!\n" % name)
        raise SystemExit(-1)


def sblas_spmm_csr_v1(pA, pB, pC, alpha, beta, n_gpu):
    """Method 1: A replicated, B/C split by columns; no inter-GPU traffic.  Like the reference
    (spmm.h:151) the host copy of each C slice is refreshed before returning."""
    assert pA.width == pB.height and pA.height == pC.height and pB.width == pC.width
    _check_colmajor(pB, pC, "SBLAS_SPMM_CSR_V1")
    print("sblas_spmm_csr_v1 ready to start")           # spmm.h:99
    assert pA.policy == M.replicate and pB.policy == M.segment and pC.policy == M.segment
    assert pB.dev_layout == pC.dev_layout
    dt = pA.dtype
    dims = _i32(pB.dim_gpu)
    ldb = np.ascontiguousarray(pB.ld_gpu, np.int64)
    ldc = np.ascontiguousarray(pC.ld_gpu, np.int64)
    fn = getattr(lib(), "sblas_mgpu_spmm_csr_v1_" + _sfx(dt))
    check(fn(C.c_uint(n_gpu), C.c_int32(pA.height), C.c_int32(pA.width), C.c_int32(pA.nnz),
             _ptr_array(pA.csrRowPtr_gpu), _ptr_array(pA.csrColIdx_gpu), _ptr_array(pA.csrVal_gpu),
             dims.ctypes.data_as(C.c_void_p), C.c_int32(pB.dev_layout), _ptr_array(pB.val_gpu),
             ldb.ctypes.data_as(C.c_void_p), _ptr_array(pC.val_gpu), ldc.ctypes.data_as(C.c_void_p),
             _sc(dt, alpha), _sc(dt, beta), _plan_array(pA.plans)), "sblas_spmm_csr_v1")
    for i in range(n_gpu):
        pC.sync2cpu(i)                                   # spmm.h:151


def sblas_spmm_csr_v2(pA, pB, pC, alpha, beta, n_gpu):
    """Method 2: A nnz-segmented, B/C replicated; every replica of C is updated."""
    assert pA.width == pB.height and pA.height == pC.height and pB.width == pC.width
    _check_colmajor(pB, pC, "SBLAS_SPMM_CSR_V2")
    assert pA.policy == M.segment and pB.policy == M.replicate and pC.policy == M.replicate
    assert pB.dev_layout == pC.dev_layout
    dt = pA.dtype
    ng, s, e = _i32(pA.nnz_gpu), _i32(pA.starting_row_gpu), _i32(pA.stoping_row_gpu)
    fn = getattr(lib(), "sblas_mgpu_spmm_csr_v2_" + _sfx(dt))
    check(fn(C.c_uint(n_gpu), C.c_int32(pA.height), C.c_int32(pA.width),
             _ptr_array(pA.csrRowPtr_gpu), _ptr_array(pA.csrColIdx_gpu), _ptr_array(pA.csrVal_gpu),
             ng.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p),
             e.ctypes.data_as(C.c_void_p), C.c_int32(pB.width), C.c_int32(pB.dev_layout),
             _ptr_array(pB.val_gpu), C.c_int64(pB.ld_gpu[0]), _ptr_array(pC.val_gpu),
             C.c_int64(pC.ld_gpu[0]), _sc(dt, alpha), _sc(dt, beta), _plan_array(pA.plans)),
          "sblas_spmm_csr_v2")
