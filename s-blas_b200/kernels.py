"""Single-device entry points and the host partitioners, on torch tensors / numpy arrays.

Every function here is a direct call through the C ABI (include/sblas_b200.h); torch only
owns the device memory and the stream.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import check, lib, make_opts, stream_ptr, vp


def _sfx(dtype):
    import torch
    if dtype in (torch.float64, np.float64) or dtype == np.dtype("float64"):
        return "f64"
    if dtype in (torch.float32, np.float32) or dtype == np.dtype("float32"):
        return "f32"
    raise TypeError("DataType must be float32 or float64 (utility.h:304-306), got %r" % (dtype,))


def _scalar(dtype, v):
    return C.c_double(float(v)) if _sfx(dtype) == "f64" else C.c_float(float(v))


def device_count():
    return int(lib().sblas_device_count())


# ---- host integer layer ------------------------------------------------------------------------
def csr_find_row(rowptr, idx):
    """utility.h:293-300."""
    rp = np.ascontiguousarray(rowptr, np.int32)
    return int(lib().sblas_csr_find_row_i32(vp(rp), C.c_int32(len(rp) - 1), C.c_int32(idx)))


def csr_partition_nnz(rowptr, n_gpu):
    """matrix.h:360-375 -> dict(avg, nnz_gpu, start_row, stop_row, rowptr=[local row pointers])."""
    rp = np.ascontiguousarray(rowptr, np.int32)
    h, nnz = len(rp) - 1, int(rp[-1])
    avg = C.c_int32()
    ng, s, e = (np.zeros(n_gpu, np.int32) for _ in range(3))
    check(lib().sblas_csr_partition_nnz_i32(vp(rp), C.c_int32(h), C.c_int32(nnz), C.c_uint(n_gpu),
                                            C.byref(avg), vp(ng), vp(s), vp(e)), "partition_nnz")
    local = []
    for i in range(n_gpu):
        if ng[i] <= 0 or s[i] < 0 or e[i] < s[i]:
            local.append(np.zeros(0, np.int32))
            continue
        out = np.zeros(int(e[i]) - int(s[i]) + 2, np.int32)
        check(lib().sblas_csr_segment_rowptr_i32(vp(rp), avg, C.c_uint(i), C.c_int32(int(s[i])),
                                                 C.c_int32(int(e[i])), C.c_int32(int(ng[i])),
                                                 vp(out)), "segment_rowptr")
        local.append(out)
    return dict(avg=int(avg.value), nnz_gpu=ng, start_row=s, stop_row=e, rowptr=local)


def dense_split(first_order, n_gpu):
    """matrix.h:557-567 -> (avg_val, dim_gpu)."""
    avg = C.c_int32()
    d = np.zeros(n_gpu, np.int32)
    check(lib().sblas_dense_split_i32(C.c_int32(first_order), C.c_uint(n_gpu), C.byref(avg), vp(d)),
          "dense_split")
    return int(avg.value), d


def csr_partition_rows(rowptr, n_gpu):
    """Row-aligned nnz-balanced panels -> row_begin[n_gpu+1]."""
    rp = np.ascontiguousarray(rowptr, np.int32)
    out = np.zeros(n_gpu + 1, np.int32)
    check(lib().sblas_csr_partition_rows_i32(vp(rp), C.c_int32(len(rp) - 1), C.c_uint(n_gpu),
                                             vp(out)), "partition_rows")
    return out


def read_mtx(path, threads=None):
    """mmio_highlevel.h:7,130 -> (h, w, rowptr, colidx, val[f64]) in file order.  `threads`: host
    threads of the parser / counting sort (None or 0 = automatic, 1 = serial); same result for any."""
    lib().sblas_mtx_set_threads(C.c_int(int(threads or 0)))
    m, n, nnz, sym = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    check(lib().sblas_mtx_info(path.encode(), C.byref(m), C.byref(n), C.byref(nnz), C.byref(sym)),
          "mtx_info")
    rp = np.zeros(m.value + 1, np.int32)
    ci = np.zeros(nnz.value, np.int32)
    v = np.zeros(nnz.value, np.float64)
    check(lib().sblas_mtx_data(path.encode(), vp(rp), vp(ci), vp(v)), "mtx_data")
    return m.value, n.value, rp, ci, v


def read_mtx_coo(path, threads=None):
    """The coordinate entries as the file holds them (CooSparseMatrix(const char*), matrix.h:142-176):
    -> (h, w, row, col, val[f64]), 0-based, file order, no symmetric expansion."""
    lib().sblas_mtx_set_threads(C.c_int(int(threads or 0)))
    m, n, nz = C.c_int32(), C.c_int32(), C.c_int32()
    check(lib().sblas_mtx_coo_info(path.encode(), C.byref(m), C.byref(n), C.byref(nz)), "mtx_coo_info")
    r = np.zeros(nz.value, np.int32)
    c = np.zeros(nz.value, np.int32)
    v = np.zeros(nz.value, np.float64)
    check(lib().sblas_mtx_coo_data(path.encode(), vp(r), vp(c), vp(v)), "mtx_coo_data")
    return m.value, n.value, r, c, v


# ---- format conversion on the device (matrix.h:40-91, :221-237) -------------------------------------
def csr2csc(m, n, rowptr, colidx, val, stream=None):
    """CsrToCsc (matrix.h:40-64) on device tensors -> (cscColPtr, cscRowIdx, cscVal)."""
    import torch
    nnz = val.numel()
    cp = torch.empty(n + 1, dtype=torch.int32, device=val.device)
    ri = torch.empty(nnz, dtype=torch.int32, device=val.device)
    cv = torch.empty(nnz, dtype=val.dtype, device=val.device)
    fn = getattr(lib(), "sblas_csr2csc_" + _sfx(val.dtype))
    check(fn(stream_ptr(stream), C.c_int32(m), C.c_int32(n), C.c_int64(nnz), vp(rowptr), vp(colidx), vp(val),
             vp(cp), vp(ri), vp(cv)), "csr2csc")
    return cp, ri, cv


def csc2csr(n, m, colptr, rowidx, val, stream=None):
    """CscToCsr (matrix.h:67-91; the reference's argument order n, m) -> (csrRowPtr, csrColIdx, csrVal)."""
    import torch
    nnz = val.numel()
    rp = torch.empty(m + 1, dtype=torch.int32, device=val.device)
    ci = torch.empty(nnz, dtype=torch.int32, device=val.device)
    cv = torch.empty(nnz, dtype=val.dtype, device=val.device)
    fn = getattr(lib(), "sblas_csc2csr_" + _sfx(val.dtype))
    check(fn(stream_ptr(stream), C.c_int32(n), C.c_int32(m), C.c_int64(nnz), vp(colptr), vp(rowidx), vp(val),
             vp(rp), vp(ci), vp(cv)), "csc2csr")
    return rp, ci, cv


def coo_sort(height, row, col, val, stream=None):
    """CooSparseMatrix::sortByRow (matrix.h:221-237), in place on device tensors, stable."""
    fn = getattr(lib(), "sblas_coo_sort_" + _sfx(val.dtype))
    check(fn(stream_ptr(stream), C.c_int64(val.numel()), C.c_int32(height), vp(row), vp(col), vp(val)), "coo_sort")
    return row, col, val


# ---- plan -------------------------------------------------------------------------------------------
class Plan:
    """Per-device CSR plan (row statistics + tile tables).  Keeps the rowptr tensor alive."""

    def __init__(self, rowptr, nnz, stream=None):
        self.rowptr = rowptr
        self.rows = rowptr.numel() - 1
        self.nnz = int(nnz)
        self.handle = C.c_void_p()
        import torch
        with torch.cuda.device(rowptr.device):
            check(lib().sblas_plan_create_i32(stream_ptr(stream), C.c_int32(self.rows),
                                              C.c_int64(self.nnz), vp(rowptr),
                                              C.byref(self.handle)), "plan_create")

    def build_rowblocks(self, colidx, val, R=0, stream=None):
        """Builds the R x 1 row-block mirror of A for SpMM (kept only when fill >= 0.5);
        returns (R, n_blocks, fill).  colidx / val must stay alive as long as the plan."""
        import torch
        fn = getattr(lib(), "sblas_plan_build_rowblocks_" + _sfx(val.dtype))
        with torch.cuda.device(val.device):
            check(fn(stream_ptr(stream), self.handle, vp(colidx), vp(val), C.c_int32(R)), "plan_build_rowblocks")
        return self.rowblock_stats()

    def prepare(self, colidx, cols, elem_size, stream=None):
        """sblas_plan_prepare_i32: col_jump statistic + the staged kernel's tile table now, so that no SpMV
        call on this plan synchronises (what CsrSparseMatrix::sync2gpu does)."""
        import torch
        with torch.cuda.device(colidx.device):
            check(lib().sblas_plan_prepare_i32(stream_ptr(stream), self.handle, vp(colidx), C.c_int32(cols),
                                               C.c_int32(elem_size)), "plan_prepare")
        return self

    def build_colpanels(self, colidx, val, cols, n_panels=0, stream=None):
        """Column-panel mirror of A for gather-heavy SpMV (sblas_plan_build_colpanels_*); n_panels = 0: only
        when it pays.  Returns (n_panels, panel_cols), (0, 0) when no mirror was built."""
        import torch
        fn = getattr(lib(), "sblas_plan_build_colpanels_" + _sfx(val.dtype))
        with torch.cuda.device(val.device):
            check(fn(stream_ptr(stream), self.handle, vp(colidx), vp(val), C.c_int32(cols), C.c_int32(n_panels)),
                  "plan_build_colpanels")
        return self.colpanel_stats()

    def colpanel_stats(self):
        n, w = C.c_int32(), C.c_int32()
        check(lib().sblas_plan_colpanel_stats(self.handle, C.byref(n), C.byref(w)), "plan_colpanel_stats")
        return int(n.value), int(w.value)

    def colpanel_format(self):
        """(is_sell, stored_entries) of the column-panel mirror (sblas_plan_colpanel_format)."""
        a, b = C.c_int32(), C.c_int64()
        check(lib().sblas_plan_colpanel_format(self.handle, C.byref(a), C.byref(b)), "plan_colpanel_format")
        return bool(a.value), int(b.value)

    def rowblock_stats(self):
        R, nb, fill = C.c_int32(), C.c_int64(), C.c_double()
        check(lib().sblas_plan_rowblock_stats(self.handle, C.byref(R), C.byref(nb), C.byref(fill)),
              "plan_rowblock_stats")
        return int(R.value), int(nb.value), float(fill.value)

    def stats(self):
        st = L.CsrStats()
        check(lib().sblas_plan_stats(self.handle, C.byref(st)), "plan_stats")
        return {f: getattr(st, f) for f, _ in st._fields_}

    def col_jump(self):
        """mean first-column jump between consecutive rows (-1 = not measured yet), see sblas_plan_col_jump"""
        out = C.c_double()
        check(lib().sblas_plan_col_jump(self.handle, C.byref(out)), "plan_col_jump")
        return out.value

    def __del__(self):
        try:
            if self.handle:
                lib().sblas_plan_destroy(self.handle)
                self.handle = C.c_void_p()
        except Exception:
            pass


def _h(plan):
    return plan.handle if plan is not None else None


# ---- kernels ----------------------------------------------------------------------------------------
def spmv(rowptr, colidx, val, x, y, alpha, beta, plan=None, opts=None, cols=None, stream=None):
    """y <- beta*y + alpha*A*x on the tensors' device (spmv.h:16-32 semantics)."""
    rows, nnz = rowptr.numel() - 1, val.numel()
    cols = x.numel() if cols is None else cols
    fn = getattr(lib(), "sblas_spmv_csr_" + _sfx(val.dtype))
    check(fn(stream_ptr(stream), C.c_int32(rows), C.c_int32(cols), C.c_int64(nnz), vp(rowptr),
             vp(colidx), vp(val), vp(x), vp(y), _scalar(val.dtype, alpha), _scalar(val.dtype, beta),
             _h(plan), C.byref(opts) if opts is not None else None), "spmv")
    return y


def spmv_noalloc(rowptr, colidx, val, x, y, alpha, beta, cols=None, stream=None):
    """sblas_spmv_csr_noalloc_f64: y <- beta*y + alpha*A*x with the scratch in a fixed device array of the library
    (no cudaMalloc, no plan); fp64, rows + nnz <= 201 326 592; overlapping calls race on the scratch."""
    rows, nnz = rowptr.numel() - 1, val.numel()
    cols = x.numel() if cols is None else cols
    check(lib().sblas_spmv_csr_noalloc_f64(stream_ptr(stream), C.c_int32(rows), C.c_int32(cols), C.c_int64(nnz),
                                           vp(rowptr), vp(colidx), vp(val), vp(x), vp(y), C.c_double(float(alpha)),
                                           C.c_double(float(beta))), "spmv_noalloc")
    return y


def spmm(rowptr, colidx, val, k, layout, B, ldb, Cm, ldc, alpha, beta, cols, plan=None, opts=None,
         stream=None):
    """C <- beta*C + alpha*A*B (spmm.h:55-68 semantics); layout ROW_MAJOR / COL_MAJOR."""
    rows, nnz = rowptr.numel() - 1, val.numel()
    fn = getattr(lib(), "sblas_spmm_csr_" + _sfx(val.dtype))
    check(fn(stream_ptr(stream), C.c_int32(rows), C.c_int32(cols), C.c_int64(nnz), vp(rowptr),
             vp(colidx), vp(val), C.c_int32(k), C.c_int32(layout), vp(B), C.c_int64(ldb), vp(Cm),
             C.c_int64(ldc), _scalar(val.dtype, alpha), _scalar(val.dtype, beta), _h(plan),
             C.byref(opts) if opts is not None else None), "spmm")
    return Cm


def set_spmm_exact(on):
    """Process-wide default arithmetic of SpMM (sblas_set_spmm_exact): True = separately rounded mul/add,
    bit-identical to spmm.h:56-68; False = fused multiply-add.  Returns the previous setting."""
    return bool(lib().sblas_set_spmm_exact(C.c_int(1 if on else 0)))


def axpby(v0, v1, alpha, beta, stream=None):
    """kernel.h:28-37: v0 = v0*beta + v1*alpha."""
    fn = getattr(lib(), "sblas_axpby_" + _sfx(v0.dtype))
    check(fn(stream_ptr(stream), vp(v0), vp(v1), _scalar(v0.dtype, alpha), _scalar(v0.dtype, beta),
             C.c_int64(v0.numel())), "axpby")
    return v0


def scal_add(v, val, beta, stream=None):
    """kernel.h:19-25: v = v*beta + val."""
    fn = getattr(lib(), "sblas_scal_add_" + _sfx(v.dtype))
    check(fn(stream_ptr(stream), vp(v), _scalar(v.dtype, val), _scalar(v.dtype, beta),
             C.c_int64(v.numel())), "scal_add")
    return v


def transpose(src, in_ld, n_major, n_minor, dst, out_ld, stream=None):
    """dst[b*out_ld + a] = src[a*in_ld + b]."""
    fn = getattr(lib(), "sblas_transpose_" + _sfx(src.dtype))
    check(fn(stream_ptr(stream), vp(src), C.c_int64(in_ld), C.c_int64(n_major), C.c_int64(n_minor),
             vp(dst), C.c_int64(out_ld)), "transpose")
    return dst


def dense_upload_rows(h, h_ld, r0, nr, k, d_dst, ld, stream=None):
    """sblas_dense_upload_rows_*: rows [r0, r0+nr) of the col-major host tensor h (pinned) -> row-major device
    block d_dst [nr x ld], streamed on `stream` (copy of chunk i+1 overlaps the conversion of chunk i)."""
    fn = getattr(lib(), "sblas_dense_upload_rows_" + _sfx(d_dst.dtype))
    check(fn(stream_ptr(stream), vp(h), C.c_int64(h_ld), C.c_int64(r0), C.c_int64(nr), C.c_int32(k), vp(d_dst),
             C.c_int64(ld)), "dense_upload_rows")
    return d_dst


def dense_download_rows(d_src, ld, nr, k, h, h_ld, r0, stream=None):
    """sblas_dense_download_rows_*: row-major device block d_src [nr x ld] -> rows [r0, r0+nr) of the col-major
    host tensor h; `stream` completes when the bytes are on the host."""
    fn = getattr(lib(), "sblas_dense_download_rows_" + _sfx(d_src.dtype))
    check(fn(stream_ptr(stream), vp(d_src), C.c_int64(ld), C.c_int64(nr), C.c_int32(k), vp(h), C.c_int64(h_ld),
             C.c_int64(r0)), "dense_download_rows")
    return h


def carry_fixup(rows, first, last, carry, k, layout, out, ld, alpha, beta, stream=None):
    rows, first, last = (np.ascontiguousarray(a, np.int32) for a in (rows, first, last))
    fn = getattr(lib(), "sblas_carry_fixup_" + _sfx(out.dtype))
    check(fn(stream_ptr(stream), C.c_int32(len(rows)), vp(rows), vp(first), vp(last), vp(carry),
             C.c_int32(k), C.c_int32(layout), vp(out), C.c_int64(ld), _scalar(out.dtype, alpha),
             _scalar(out.dtype, beta)), "carry_fixup")
    return out
