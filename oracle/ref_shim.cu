// oracle/ref_shim.cu — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" window onto the UNMODIFIED reference (tartarughina/S-BLAS) CPU
// verification path, compiled from the reference's own headers where they lie
// (-I/root/reference; nothing is copied into this repo).  Built by
// oracle/Makefile into oracle/_ref/libsblas_ref.so (git-ignored, travels to the
// GPU box).  It is used (1) to pin the C restatement in oracle/sblas_oracle.c,
// (2) to generate tests/golden/*, (3) as the "reference" CPU baseline in
// bench.py (--impl reference and the cpu_baseline leg).
//
// How the reference objects are fed without a GPU: CsrSparseMatrix / DenseMatrix /
// DenseVector have public pointer fields and default constructors that allocate
// nothing (matrix.h:278-288, :512-517, :655-658).  We point those fields at the
// caller's buffers, call the reference's template functions, and null the fields
// again before the destructors run (the destructors only free non-NULL pointers,
// utility.h:86-110), so no cudaMallocHost / driver is ever needed.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <algorithm>
using std::min;
using std::max;
#include "sblas.h"   // the reference's umbrella header (sblas.h:18-19)

namespace {
template <typename T>
struct CsrView {
  CsrSparseMatrix<int, T> m;
  CsrView(int h, int w, int nnz, const int* rp, const int* ci, const T* v) {
    m.height = h; m.width = w; m.nnz = nnz;
    m.csrRowPtr = const_cast<int*>(rp);
    m.csrColIdx = const_cast<int*>(ci);
    m.csrVal = const_cast<T*>(v);
  }
  ~CsrView() { m.csrRowPtr = NULL; m.csrColIdx = NULL; m.csrVal = NULL; m.n_gpu = 0; }
};
template <typename T>
struct VecView {
  DenseVector<int, T> v;
  VecView(int len, T* p) { v.length = len; v.val = p; v.val_gpu = NULL; v.n_gpu = 0; v.policy = none; }
  ~VecView() { v.val = NULL; v.val_gpu = NULL; v.n_gpu = 0; }
};
template <typename T>
struct MatView {
  DenseMatrix<int, T> m;
  MatView(int h, int w, T* p, int order) {
    m.height = h; m.width = w; m.val = p; m.order = order ? col_major : row_major;
  }
  ~MatView() { m.val = NULL; m.val_gpu = NULL; m.dim_gpu = NULL; m.n_gpu = 0; }
};
}  // namespace

extern "C" {

// spmv.h:16-32 (requires square A: spmv.h:21)
void ref_spmv_csr_cpu_f64(int h, int w, int nnz, const int* rp, const int* ci, const double* v,
                          const double* x, double* y, double alpha, double beta) {
  CsrView<double> A(h, w, nnz, rp, ci, v);
  VecView<double> B(w, const_cast<double*>(x)), C(h, y);
  sblas_spmv_csr_cpu<int, double>(&A.m, &B.v, &C.v, alpha, beta);
}
void ref_spmv_csr_cpu_f32(int h, int w, int nnz, const int* rp, const int* ci, const float* v,
                          const float* x, float* y, float alpha, float beta) {
  CsrView<float> A(h, w, nnz, rp, ci, v);
  VecView<float> B(w, const_cast<float*>(x)), C(h, y);
  sblas_spmv_csr_cpu<int, float>(&A.m, &B.v, &C.v, alpha, beta);
}
// spmm.h:30-70; B and C column-major (order=1) is the only exercised branch.
void ref_spmm_csr_cpu_f64(int h, int w, int nnz, const int* rp, const int* ci, const double* v,
                          int k, const double* B, double* C, double alpha, double beta) {
  CsrView<double> A(h, w, nnz, rp, ci, v);
  MatView<double> mB(w, k, const_cast<double*>(B), 1), mC(h, k, C, 1);
  sblas_spmm_csr_cpu<int, double>(&A.m, &mB.m, &mC.m, alpha, beta);
}
void ref_spmm_csr_cpu_f32(int h, int w, int nnz, const int* rp, const int* ci, const float* v,
                          int k, const float* B, float* C, float alpha, float beta) {
  CsrView<float> A(h, w, nnz, rp, ci, v);
  MatView<float> mB(w, k, const_cast<float*>(B), 1), mC(h, k, C, 1);
  sblas_spmm_csr_cpu<int, float>(&A.m, &mB.m, &mC.m, alpha, beta);
}

// utility.h:293-300
int ref_find_row(const int* rp, int h, int nnz_idx) {
  return csr_findRowIdxUsingNnzIdx<int>(rp, h, nnz_idx);
}

// The partition arithmetic of CsrSparseMatrix::sync2gpu(segment), matrix.h:360-375,
// re-evaluated expression by expression with the reference's own
// csr_findRowIdxUsingNnzIdx (sync2gpu itself needs cudaMalloc, so it cannot run here).
// rowptr_out receives the n_gpu local row-pointer arrays back to back;
// rowptr_off[i] is the start of GPU i's array, rowptr_off[n_gpu] the total.
int ref_partition_nnz(const int* csrRowPtr, int height, int nnz, unsigned n_gpu, int* avg_out,
                      int* nnz_gpu, int* starting_row_gpu, int* stoping_row_gpu, int* rowptr_out,
                      long* rowptr_off) {
  int avg_nnz = ceil((float)nnz / n_gpu);  // matrix.h:360 (float32 on purpose)
  *avg_out = avg_nnz;
  long off = 0;
  for (unsigned i = 0; i < n_gpu; i++) {
    int row_starting_nnz = i * avg_nnz;                    // matrix.h:363
    int row_stoping_nnz = min((int)((i + 1) * avg_nnz), nnz) - 1;  // matrix.h:364
    nnz_gpu[i] = row_stoping_nnz - row_starting_nnz + 1;   // matrix.h:365
    starting_row_gpu[i] = csr_findRowIdxUsingNnzIdx(csrRowPtr, height, row_starting_nnz);
    stoping_row_gpu[i] = csr_findRowIdxUsingNnzIdx(csrRowPtr, height, row_stoping_nnz);
    long num = (long)stoping_row_gpu[i] - starting_row_gpu[i] + 2;  // matrix.h:401
    rowptr_off[i] = off;
    if (rowptr_out) {
      int* fixedRowPtr = rowptr_out + off;
      fixedRowPtr[0] = 0;                                   // matrix.h:372
      for (int k = 1; k < num - 1; k++)
        fixedRowPtr[k] = csrRowPtr[starting_row_gpu[i] + k] - i * avg_nnz;  // matrix.h:374
      fixedRowPtr[num - 1] = nnz_gpu[i];                    // matrix.h:375
    }
    off += num;
  }
  rowptr_off[n_gpu] = off;
  return 0;
}

// DenseMatrix::sync2gpu(segment) split arithmetic, matrix.h:557-567.
void ref_dense_split(int first_order, unsigned n_gpu, int* avg_out, int* dim_gpu) {
  int avg_val = ceil((double)first_order / n_gpu);
  *avg_out = avg_val;
  for (unsigned i = 0; i < n_gpu; i++)
    dim_gpu[i] = min((int)((i + 1) * avg_val), first_order) - i * avg_val;
}

// Dense random constructor fill, matrix.h:522-524 / utility.h:197 (srand(211), glibc rand()).
void ref_rand_fill_f64(double* p, long n) {
  srand(RAND_INIT_SEED);
  for (long i = 0; i < n; i++) p[i] = (double)rand0to1();
}
void ref_rand_fill_f32(float* p, long n) {
  srand(RAND_INIT_SEED);
  for (long i = 0; i < n; i++) p[i] = (float)rand0to1();
}

// .mtx -> CSR exactly as CsrSparseMatrix(const char*) does it (matrix.h:300-318):
// mmio_info then mmio_data (mmio_highlevel.h:7, :130).
int ref_mtx_info(const char* fn, int* m, int* n, int* nnz, int* sym) {
  return mmio_info(m, n, nnz, sym, fn);
}
int ref_mtx_data(const char* fn, int* rp, int* ci, double* v) { return mmio_data(rp, ci, v, fn); }

// matrix.h:40-64 and :67-91, the reference's own templates (plain CPU code)
void ref_csr2csc_f64(int m, int n, int nnz, const int* rp, const int* ci, const double* v, int* cscRowIdx,
                     int* cscColPtr, double* cscVal) {
  CsrToCsc<int, double>(m, n, nnz, rp, ci, v, cscRowIdx, cscColPtr, cscVal);
}
void ref_csr2csc_f32(int m, int n, int nnz, const int* rp, const int* ci, const float* v, int* cscRowIdx,
                     int* cscColPtr, float* cscVal) {
  CsrToCsc<int, float>(m, n, nnz, rp, ci, v, cscRowIdx, cscColPtr, cscVal);
}
void ref_csc2csr_f64(int n, int m, int nnz, const int* cp, const int* ri, const double* v, int* csrColIdx,
                     int* csrRowPtr, double* csrVal) {
  CscToCsr<int, double>(n, m, nnz, cp, ri, v, csrColIdx, csrRowPtr, csrVal);
}

// utility.h:182-191
int ref_check_equal_f64(const double* x, const double* y, size_t m) { return check_equal<double>(x, y, m); }
int ref_check_equal_f32(const float* x, const float* y, size_t m) { return check_equal<float>(x, y, m); }

}  // extern "C"
