"""Host-side mirror of the reference's data model for the hot path (matrix.h):
CsrSparseMatrix (:276-453), DenseMatrix (:510-649), DenseVector (:653-740), the
GpuSharePolicy / MajorOrder enums (:34,:37).  Same field names, method names and
partition arithmetic; host arrays are numpy (pinned through torch when a GPU is present),
per-GPU arrays are torch CUDA tensors (torch = device-memory plumbing only).

Single process driving GPUs 0..n_gpu-1, like the reference (spmv.h:52-55).  The
one-process-per-GPU path used by bench.py at N > 1 lives in dist.py.
"""
import os

import numpy as np

from . import kernels as K
from ._lib import COL_MAJOR, ROW_MAJOR

# enum GpuSharePolicy (matrix.h:34) / enum MajorOrder (matrix.h:37)
none, replicate, segment = 0, 1, 2
row_major, col_major = 0, 1


def _torch():
    import torch
    return torch


def _np_dtype(dtype):
    return np.dtype(dtype)


def _tdtype(npdt):
    torch = _torch()
    return torch.float64 if np.dtype(npdt) == np.float64 else torch.float32


def physical(i):
    """CUDA device ordinal of logical GPU i (SBLAS_DEVICE_MAP, see sblas_device_of)."""
    from ._lib import lib
    return int(lib().sblas_device_of(int(i)))


def _dev(i):
    return _torch().device("cuda", physical(i))


def _to_dev(a, i):
    torch = _torch()
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(_dev(i), non_blocking=False)


class CsrSparseMatrix:
    """matrix.h:276-453.  Construct from a .mtx path (matrix.h:300-330) or from host arrays."""

    def __init__(self, path=None, *, height=None, width=None, rowptr=None, colidx=None, val=None,
                 dtype=np.float64):
        self.csrRowPtr_gpu = self.csrColIdx_gpu = self.csrVal_gpu = None
        self.nnz_gpu = self.starting_row_gpu = self.stoping_row_gpu = None
        self.plans = None
        self.n_gpu, self.policy = 0, none
        if path is not None:
            h, w, rp, ci, v = K.read_mtx(path)
            print("input matrix A: ( %i, %i ) nnz = %i" % (h, w, len(ci)))   # matrix.h:318
            self.height, self.width = h, w
            self.csrRowPtr, self.csrColIdx = rp, ci
            self.csrVal = v.astype(_np_dtype(dtype))
        else:
            self.height, self.width = int(height), int(width)
            self.csrRowPtr = np.ascontiguousarray(rowptr, np.int32)
            self.csrColIdx = np.ascontiguousarray(colidx, np.int32)
            self.csrVal = np.ascontiguousarray(val)
        self.nnz = int(self.csrRowPtr[-1])
        self.dtype = self.csrVal.dtype

    def sync2gpu(self, n_gpu, policy, quiet=False):
        """matrix.h:331-397: replicate, or nnz-segment (bit-exact partition tables)."""
        assert n_gpu != 0 and policy != none
        self.n_gpu, self.policy = n_gpu, policy
        torch = _torch()
        self.csrRowPtr_gpu, self.csrColIdx_gpu, self.csrVal_gpu, self.plans = [], [], [], []
        if policy == replicate:
            for i in range(n_gpu):
                self.csrRowPtr_gpu.append(_to_dev(self.csrRowPtr, i))
                self.csrColIdx_gpu.append(_to_dev(self.csrColIdx, i))
                self.csrVal_gpu.append(_to_dev(self.csrVal, i))
                with torch.cuda.device(physical(i)):
                    self.plans.append(K.Plan(self.csrRowPtr_gpu[i], self.nnz))
                    self.plans[-1].build_rowblocks(self.csrColIdx_gpu[i], self.csrVal_gpu[i])
        else:
            p = K.csr_partition_nnz(self.csrRowPtr, n_gpu)
            self.avg_nnz = p["avg"]
            self.nnz_gpu, self.starting_row_gpu, self.stoping_row_gpu = (
                p["nnz_gpu"], p["start_row"], p["stop_row"])
            for i in range(n_gpu):
                lo = i * p["avg"]
                n_i = max(int(self.nnz_gpu[i]), 0)
                self.csrRowPtr_gpu.append(_to_dev(p["rowptr"][i], i))
                self.csrColIdx_gpu.append(_to_dev(self.csrColIdx[lo:lo + n_i], i))
                self.csrVal_gpu.append(_to_dev(self.csrVal[lo:lo + n_i], i))
                with torch.cuda.device(physical(i)):
                    self.plans.append(K.Plan(self.csrRowPtr_gpu[i], n_i) if n_i > 0 and
                                      len(p["rowptr"][i]) else None)
                    if self.plans[-1] is not None:
                        self.plans[-1].build_rowblocks(self.csrColIdx_gpu[i], self.csrVal_gpu[i])
                if not quiet:                                               # matrix.h:388
                    print("gpu-%d,start-row:%d,stop-row:%d,num-rows:%d,num-nnz:%d" % (
                        i, self.starting_row_gpu[i], self.stoping_row_gpu[i],
                        self.get_gpu_row_ptr_num(i), self.nnz_gpu[i]))

    def get_gpu_row_ptr_num(self, i):                                       # matrix.h:398-404
        if self.nnz_gpu is None:
            return 0
        return int(self.stoping_row_gpu[i]) - int(self.starting_row_gpu[i]) + 2


class DenseVector:
    """matrix.h:653-740.  DenseVector(n) = glibc-random ctor is in tests (needs srand/rand)."""

    def __init__(self, length, val=None, dtype=np.float64, data=None):
        self.length = int(length)
        if data is not None:
            self.val = np.ascontiguousarray(data, _np_dtype(dtype)).copy()
        else:
            self.val = np.full(self.length, 0.0 if val is None else val, _np_dtype(dtype))
        self.val_gpu, self.n_gpu, self.policy = None, 0, none

    def sync2gpu(self, n_gpu, policy):                                      # matrix.h:688-704
        assert n_gpu != 0 and policy != none
        assert policy != segment, "vectors are always replicated (matrix.h:693)"
        self.n_gpu, self.policy = n_gpu, policy
        self.val_gpu = [_to_dev(self.val, i) for i in range(n_gpu)]

    def sync2cpu(self, i_gpu):                                              # matrix.h:705-713
        assert i_gpu < self.n_gpu and self.val_gpu is not None
        self.val[:] = self.val_gpu[i_gpu].cpu().numpy()

    def plusDenseVectorGPU(self, dv, alpha, beta):                          # matrix.h:714-726
        torch = _torch()
        for i in range(self.n_gpu):
            with torch.cuda.device(physical(i)):
                K.axpby(self.val_gpu[i], dv.val_gpu[i], alpha, beta)
        for i in range(self.n_gpu):
            torch.cuda.synchronize(physical(i))

    def get_vec_length(self):
        return self.length


class DenseMatrix:
    """matrix.h:510-649.  Host storage in `order` (col_major for every SpMM operand,
    spmm.h:91-98).  The DEVICE copy of a col_major matrix is kept row-major (leading dimension
    padded to 16 bytes) unless device_layout="col": sync2gpu / sync2cpu transpose on the
    device, so callers that only use this class see the reference's behaviour."""

    def __init__(self, height, width, val=None, order=col_major, dtype=np.float64, data=None):
        self.height, self.width, self.order = int(height), int(width), order
        n = self.height * self.width
        if data is not None:
            self.val = np.ascontiguousarray(data, _np_dtype(dtype)).reshape(-1).copy()
            assert self.val.size == n
        else:
            self.val = np.full(n, 0.0 if val is None else val, _np_dtype(dtype))
        self.val_gpu, self.dim_gpu, self.n_gpu, self.policy = None, None, 0, none
        self.dev_layout, self.ld_gpu = None, None

    def get_mtx_num(self):
        return self.height * self.width

    def _second_order(self):
        return self.width if self.order == row_major else self.height

    def sync2gpu(self, n_gpu, policy, device_layout=None):                  # matrix.h:541-570
        assert n_gpu != 0 and policy != none
        torch = _torch()
        self.n_gpu, self.policy = n_gpu, policy
        lay = device_layout or os.environ.get("SBLAS_DENSE_DEVICE_LAYOUT", "row")
        # only col_major host matrices are re-laid out; row_major ones are already row-major
        self.dev_layout = ROW_MAJOR if (lay == "row" or self.order == row_major) else COL_MAJOR
        first = self.height if self.order == row_major else self.width
        second = self._second_order()
        if policy == replicate:
            self.avg_val, dims = first, [first] * n_gpu
            offs = [0] * n_gpu
        else:
            self.avg_val, d = K.dense_split(first, n_gpu)
            self.dim_gpu = d
            dims = [int(x) for x in d]
            offs = [i * self.avg_val for i in range(n_gpu)]
        self.val_gpu, self.ld_gpu = [], []
        vec = 16 // self.val.dtype.itemsize
        for i in range(n_gpu):
            if dims[i] <= 0:
                self.val_gpu.append(torch.empty(0, dtype=_tdtype(self.val.dtype), device=_dev(i)))
                self.ld_gpu.append(1)
                continue
            host = self.val[offs[i] * second:(offs[i] + dims[i]) * second]
            stage = _to_dev(host, i)          # [dims x second], `second` contiguous
            if self.order == col_major and self.dev_layout == ROW_MAJOR:
                ld = (dims[i] + vec - 1) // vec * vec
                dst = torch.zeros(second * ld, dtype=stage.dtype, device=_dev(i))
                with torch.cuda.device(physical(i)):
                    K.transpose(stage, second, dims[i], second, dst, ld)
                    torch.cuda.synchronize()
                self.val_gpu.append(dst)
                self.ld_gpu.append(ld)
            else:
                self.val_gpu.append(stage)
                self.ld_gpu.append(second)

    def sync2cpu(self, i_gpu):                                              # matrix.h:595-612
        assert self.val_gpu is not None and i_gpu < self.n_gpu
        torch = _torch()
        second = self._second_order()
        if self.policy == segment:
            dim, off = int(self.dim_gpu[i_gpu]), i_gpu * self.avg_val
        else:
            dim, off = (self.height if self.order == row_major else self.width), 0
        if dim <= 0:
            return
        src = self.val_gpu[i_gpu]
        if self.order == col_major and self.dev_layout == ROW_MAJOR:
            tmp = torch.empty(dim * second, dtype=src.dtype, device=src.device)
            with torch.cuda.device(physical(i_gpu)):
                K.transpose(src, self.ld_gpu[i_gpu], second, dim, tmp, second)
                torch.cuda.synchronize()
            src = tmp
        self.val[off * second:(off + dim) * second] = src.cpu().numpy()

    def plusDenseMatrixGPU(self, dm, alpha, beta):                          # matrix.h:613-625
        torch = _torch()
        for i in range(self.n_gpu):
            with torch.cuda.device(physical(i)):
                K.axpby(self.val_gpu[i], dm.val_gpu[i], alpha, beta)
        for i in range(self.n_gpu):
            torch.cuda.synchronize(physical(i))


# ---- format containers beside CSR (matrix.h:40-272, :457-506) -----------------------
def CsrToCsc(m, n, nnz, csrRowPtr, csrColIdx, csrVal):
    """matrix.h:40-64 on host arrays (the work runs on GPU 0): -> (cscRowIdx, cscColPtr, cscVal),
    inside a column the entries keep CSR order, bit-identical to the reference's CPU loop."""
    torch = _torch()
    with torch.cuda.device(physical(0)):
        cp, ri, cv = K.csr2csc(int(m), int(n), _to_dev(np.asarray(csrRowPtr, np.int32), 0),
                               _to_dev(np.asarray(csrColIdx, np.int32)[:nnz], 0), _to_dev(np.asarray(csrVal)[:nnz], 0))
        torch.cuda.synchronize()
    return ri.cpu().numpy(), cp.cpu().numpy(), cv.cpu().numpy()


def CscToCsr(n, m, nnz, cscColPtr, cscRowIdx, cscVal):
    """matrix.h:67-91 (the reference's argument order n, m): -> (csrColIdx, csrRowPtr, csrVal)."""
    torch = _torch()
    with torch.cuda.device(physical(0)):
        rp, ci, cv = K.csc2csr(int(n), int(m), _to_dev(np.asarray(cscColPtr, np.int32), 0),
                               _to_dev(np.asarray(cscRowIdx, np.int32)[:nnz], 0), _to_dev(np.asarray(cscVal)[:nnz], 0))
        torch.cuda.synchronize()
    return ci.cpu().numpy(), rp.cpu().numpy(), cv.cpu().numpy()


class CooSparseMatrix:
    """matrix.h:117-272: entries as the .mtx file holds them (no symmetric expansion, :165-173), sorted by
    (row, col) on GPU 0 (sortByRow, :221-237; equal pairs keep file order), then replicated or cut into
    floor(nnz / n_gpu)-sized pieces (:196-216)."""

    def __init__(self, path=None, n_gpu=0, policy=none, dtype=np.float64):
        self.cooRowIdx = self.cooColIdx = self.cooVal = None
        self.cooRowIdx_gpu = self.cooColIdx_gpu = self.cooVal_gpu = self.nnz_gpu = None
        self.nnz = self.height = self.width = 0
        self.n_gpu, self.policy = 0, none
        if path is None:
            return
        print("Loading input matrix from '%s'." % path)
        self.height, self.width, r, c, v = K.read_mtx_coo(path)
        self.nnz = len(r)
        print("Height: %d Width: %d nnz: %d" % (self.height, self.width, self.nnz))
        self.cooRowIdx, self.cooColIdx, self.cooVal = r, c, v.astype(_np_dtype(dtype))
        self.sortByRow()
        self.n_gpu, self.policy = n_gpu, policy
        if n_gpu != 0 and policy != none:
            avg = self.nnz // n_gpu                                   # matrix.h:198
            self.nnz_gpu = np.array([(min((i + 1) * avg, self.nnz) - i * avg) if policy == segment else self.nnz
                                     for i in range(n_gpu)], np.int32)
            self.cooRowIdx_gpu, self.cooColIdx_gpu, self.cooVal_gpu = [], [], []
            for i in range(n_gpu):
                lo = i * avg if policy == segment else 0
                hi = lo + int(self.nnz_gpu[i])
                self.cooRowIdx_gpu.append(_to_dev(self.cooRowIdx[lo:hi], i))
                self.cooColIdx_gpu.append(_to_dev(self.cooColIdx[lo:hi], i))
                self.cooVal_gpu.append(_to_dev(self.cooVal[lo:hi], i))

    def sortByRow(self):
        if self.nnz == 0:
            return
        torch = _torch()
        with torch.cuda.device(physical(0)):
            r, c, v = (_to_dev(a, 0) for a in (self.cooRowIdx, self.cooColIdx, self.cooVal))
            K.coo_sort(self.height, r, c, v)
            torch.cuda.synchronize()
        self.cooRowIdx, self.cooColIdx, self.cooVal = r.cpu().numpy(), c.cpu().numpy(), v.cpu().numpy()


class CscSparseMatrix:
    """matrix.h:457-506: built from a CsrSparseMatrix by CsrToCsc."""

    def __init__(self, csr=None):
        self.cscRowIdx = self.cscColPtr = self.cscVal = None
        self.nnz = self.height = self.width = 0
        if csr is None:
            return
        self.height, self.width, self.nnz = csr.height, csr.width, csr.nnz
        print("Building csc matrix from a csr matrix.")
        print("Height: %d Width: %d nnz: %d" % (self.height, self.width, self.nnz))
        self.cscRowIdx, self.cscColPtr, self.cscVal = CsrToCsc(self.height, self.width, self.nnz, csr.csrRowPtr,
                                                               csr.csrColIdx, csr.csrVal)
