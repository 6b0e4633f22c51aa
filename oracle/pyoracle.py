"""oracle/pyoracle.py — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy/ctypes front-end for the two CPU checkers:

* ``Oracle()``   -> oracle/_build/libsblas_oracle.so, the plain-C restatement
                    (oracle/sblas_oracle.c) of the reference's CPU verification path;
* ``Reference()``-> oracle/_ref/libsblas_ref.so, the UNMODIFIED reference compiled from
                    /root/reference by oracle/Makefile (present when that build ran).

Both expose the same method names so tests can run one against the other.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libsblas_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libsblas_ref.so")


def build(force=False):
    """Compile the C restatement (and the reference shim when /root/reference exists)."""
    if force or not os.path.exists(ORACLE_SO) or (
            os.path.getmtime(ORACLE_SO) < os.path.getmtime(os.path.join(HERE, "sblas_oracle.c"))):
        subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    return ORACLE_SO


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _ct(dt):
    return C.c_double if np.dtype(dt) == np.float64 else C.c_float


def _sfx(dt):
    return "f64" if np.dtype(dt) == np.float64 else "f32"


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class _Base:
    prefix = ""

    def __init__(self, path):
        self.lib = C.CDLL(path)
        self.path = path

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)


class Oracle(_Base):
    """The C restatement.  Semantics per function: see oracle/sblas_oracle.c."""
    prefix = "oracle_"

    def __init__(self):
        build()
        super().__init__(ORACLE_SO)
        self.lib.oracle_max_threads.restype = C.c_int

    kind = "port"

    def max_threads(self):
        return int(self.lib.oracle_max_threads())

    # y <- beta*y + alpha*A*x   (spmv.h:16-32); returns the updated y (in place)
    def spmv(self, rp, ci, v, x, y, alpha, beta):
        rp, ci = _i32(rp), _i32(ci)
        dt = v.dtype
        ct = _ct(dt)
        self._f("spmv_csr_" + _sfx(dt))(C.c_int(len(rp) - 1), _p(rp), _p(ci), _p(v), _p(x), _p(y),
                                        ct(alpha), ct(beta))
        return y

    def spmv_rows_omp(self, r0, r1, rp, ci, v, x, y, alpha, beta, nthreads):
        dt = v.dtype
        ct = _ct(dt)
        self._f("spmv_csr_rows_omp_" + _sfx(dt))(C.c_int(r0), C.c_int(r1), _p(rp), _p(ci), _p(v),
                                                 _p(x), _p(y), ct(alpha), ct(beta),
                                                 C.c_int(nthreads))
        return y

    # C <- beta*C + alpha*A*B, B [w x k] and C [h x k] column-major 1-D arrays (spmm.h:55-68)
    def spmm(self, rp, ci, v, k, B, Cm, alpha, beta, Bh=None, Ch=None):
        rp, ci = _i32(rp), _i32(ci)
        h = len(rp) - 1
        dt = v.dtype
        ct = _ct(dt)
        Bh = B.size // k if Bh is None else Bh
        Ch = h if Ch is None else Ch
        self._f("spmm_csr_" + _sfx(dt))(C.c_int(h), _p(rp), _p(ci), _p(v), C.c_int(k), _p(B),
                                        C.c_long(Bh), _p(Cm), C.c_long(Ch), ct(alpha), ct(beta))
        return Cm

    def spmm_rows_omp(self, r0, r1, rp, ci, v, k, B, Bh, Cm, Ch, alpha, beta, nthreads):
        dt = v.dtype
        ct = _ct(dt)
        self._f("spmm_csr_rows_omp_" + _sfx(dt))(C.c_int(r0), C.c_int(r1), _p(rp), _p(ci), _p(v),
                                                 C.c_int(k), _p(B), C.c_long(Bh), _p(Cm),
                                                 C.c_long(Ch), ct(alpha), ct(beta),
                                                 C.c_int(nthreads))
        return Cm

    def absrow_spmv(self, rp, ci, v, x):
        out = np.zeros(len(rp) - 1, np.float64)
        self.lib.oracle_absrow_spmv_f64(C.c_int(len(rp) - 1), _p(_i32(rp)), _p(_i32(ci)),
                                        _p(np.ascontiguousarray(v, np.float64)),
                                        _p(np.ascontiguousarray(x, np.float64)), _p(out))
        return out

    def axpby(self, v0, v1, alpha, beta):
        ct = _ct(v0.dtype)
        self._f("axpby_" + _sfx(v0.dtype))(_p(v0), _p(v1), ct(alpha), ct(beta), C.c_long(v0.size))
        return v0

    def check_equal(self, x, y):
        return bool(self._f("check_equal_" + _sfx(x.dtype))(_p(x), _p(y), C.c_size_t(x.size)))

    def find_row(self, rp, idx):
        rp = _i32(rp)
        return int(self._f("find_row")(_p(rp), C.c_int(len(rp) - 1), C.c_int(idx)))

    def partition_nnz(self, rp, n_gpu, float_ceil=True):
        """matrix.h:360-375.  Returns dict(avg, nnz_gpu, start_row, stop_row, rowptr=[...])."""
        rp = _i32(rp)
        h = len(rp) - 1
        nnz = int(rp[-1])
        avg = C.c_int()
        ng = np.zeros(n_gpu, np.int32)
        s = np.zeros(n_gpu, np.int32)
        e = np.zeros(n_gpu, np.int32)
        off = np.zeros(n_gpu + 1, np.int64)
        args = [_p(rp), C.c_int(h), C.c_int(nnz), C.c_uint(n_gpu)]
        if self.prefix == "oracle_":
            args.append(C.c_int(1 if float_ceil else 0))
        self._f("partition_nnz")(*args, C.byref(avg), _p(ng), _p(s), _p(e), None, _p(off))
        out = np.zeros(int(off[-1]), np.int32)
        self._f("partition_nnz")(*args, C.byref(avg), _p(ng), _p(s), _p(e), _p(out), _p(off))
        return dict(avg=int(avg.value), nnz_gpu=ng, start_row=s, stop_row=e,
                    rowptr=[out[off[i]:off[i + 1]].copy() for i in range(n_gpu)])

    def dense_split(self, first_order, n_gpu):
        avg = C.c_int()
        d = np.zeros(n_gpu, np.int32)
        self._f("dense_split")(C.c_int(first_order), C.c_uint(n_gpu), C.byref(avg), _p(d))
        return int(avg.value), d

    def rand_fill(self, n, dtype=np.float64):
        a = np.zeros(n, dtype)
        self._f("rand_fill_" + _sfx(dtype))(_p(a), C.c_long(n))
        return a

    def read_mtx(self, path):
        """-> (h, w, rowptr int32, colidx int32, val float64), mmio_highlevel.h:7,130."""
        m, n, nnz, sym = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        rc = self._f("mtx_info")(path.encode(), C.byref(m), C.byref(n), C.byref(nnz), C.byref(sym))
        if rc != 0:
            raise IOError("mtx_info(%s) -> %d" % (path, rc))
        rp = np.zeros(m.value + 1, np.int32)
        ci = np.zeros(max(nnz.value, 1), np.int32)[:nnz.value]
        v = np.zeros(max(nnz.value, 1), np.float64)[:nnz.value]
        rc = self._f("mtx_data")(path.encode(), _p(rp), _p(ci), _p(v))
        if rc != 0:
            raise IOError("mtx_data(%s) -> %d" % (path, rc))
        return m.value, n.value, rp, ci, v


    def csr2csc(self, m, n, rp, ci, v):
        """matrix.h:40-64 -> (cscColPtr, cscRowIdx, cscVal)"""
        rp, ci = _i32(rp), _i32(ci)
        nnz = len(ci)
        cp = np.zeros(n + 1, np.int32)
        ri = np.zeros(max(nnz, 1), np.int32)[:nnz]
        cv = np.zeros(max(nnz, 1), v.dtype)[:nnz]
        self._f("csr2csc_" + _sfx(v.dtype))(C.c_int(m), C.c_int(n), C.c_int(nnz), _p(rp), _p(ci), _p(v), _p(ri),
                                            _p(cp), _p(cv))
        return cp, ri, cv

    def csc2csr(self, n, m, cp, ri, v):
        """matrix.h:67-91 -> (csrRowPtr, csrColIdx, csrVal); the same loop with the roles exchanged"""
        if self.prefix == "ref_":
            assert v.dtype == np.float64
            cp, ri = _i32(cp), _i32(ri)
            nnz = len(ri)
            rp = np.zeros(m + 1, np.int32)
            ci = np.zeros(max(nnz, 1), np.int32)[:nnz]
            cv = np.zeros(max(nnz, 1), v.dtype)[:nnz]
            self._f("csc2csr_f64")(C.c_int(n), C.c_int(m), C.c_int(nnz), _p(cp), _p(ri), _p(v), _p(ci), _p(rp), _p(cv))
            return rp, ci, cv
        return self.csr2csc(n, m, cp, ri, v)

    def coo_sort(self, r, c, v):
        """matrix.h:221-237 (stable restatement) -> sorted copies"""
        r, c, v = _i32(r).copy(), _i32(c).copy(), np.ascontiguousarray(v, np.float64).copy()
        self._f("coo_sort_f64")(C.c_long(len(r)), _p(r), _p(c), _p(v))
        return r, c, v


class Reference(Oracle):
    """The reference's own code (oracle/ref_shim.cu).  Same methods; differences:
    spmv needs square A (spmv.h:21); spmm only with Bh=w, Ch=h; no *_rows_omp."""
    prefix = "ref_"
    kind = "reference"

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        _Base.__init__(self, REF_SO)

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def max_threads(self):
        return 1

    def spmv(self, rp, ci, v, x, y, alpha, beta):
        rp, ci = _i32(rp), _i32(ci)
        h = len(rp) - 1
        assert len(x) == h == len(y), "reference CPU SpMV asserts square A (spmv.h:20-21)"
        ct = _ct(v.dtype)
        self._f("spmv_csr_cpu_" + _sfx(v.dtype))(C.c_int(h), C.c_int(len(x)), C.c_int(len(ci)),
                                                 _p(rp), _p(ci), _p(v), _p(x), _p(y), ct(alpha),
                                                 ct(beta))
        return y

    def spmm(self, rp, ci, v, k, B, Cm, alpha, beta, Bh=None, Ch=None):
        rp, ci = _i32(rp), _i32(ci)
        h = len(rp) - 1
        w = B.size // k
        assert Bh in (None, w) and Ch in (None, h)
        ct = _ct(v.dtype)
        self._f("spmm_csr_cpu_" + _sfx(v.dtype))(C.c_int(h), C.c_int(w), C.c_int(len(ci)), _p(rp),
                                                 _p(ci), _p(v), C.c_int(k), _p(B), _p(Cm),
                                                 ct(alpha), ct(beta))
        return Cm

    def spmv_rows_omp(self, *a, **k):
        raise NotImplementedError("the reference CPU path is single-threaded (spmv.h:22-31)")

    spmm_rows_omp = spmv_rows_omp

    def absrow_spmv(self, *a, **k):
        raise NotImplementedError

    def axpby(self, *a, **k):
        raise NotImplementedError("kernel.h:28 is a __global__ kernel; no CPU form in the reference")
