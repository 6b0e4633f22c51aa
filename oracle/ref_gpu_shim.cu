// oracle/ref_gpu_shim.cu — TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// a "reference-GPU" comparison lane.  The UNMODIFIED reference
// (tartarughina/S-BLAS) multi-GPU path — cuSPARSE SpMV/SpMM per GPU + ncclAllReduce + axpby,
// spmv.h:36-139, spmm.h:84-161, spmm.h:164-284 — compiled from its own headers where they lie
// (-I/root/reference; nothing is copied) into oracle/_ref/libsblas_refgpu.so by oracle/Makefile
// (target `refgpu`, arch compute_100 instead of the reference Makefile's compute_80).  Only
// tools/ref_gpu_lane.py loads it: it times the library path on the SAME synthetic operands as
// bench.py and checks its output against ours.  No product path links or loads this file.
//
// The reference objects are fed exactly as in ref_shim.cu: default-constructed, public host
// fields pointed at the caller's (pageable) buffers, the reference's own sync2gpu() builds the
// device copies, and the host fields are nulled again before the destructors run so that only
// what the reference allocated is freed (utility.h:86-110).
//
// The reference prints per-GPU lines to stdout inside the ops (spmv.h:125, matrix.h:389); the
// caller redirects fd 1 around these calls.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <chrono>
using std::min;
using std::max;
#include "sblas.h"   // the reference's umbrella header

namespace {
template <typename T>
struct CsrHold {
  CsrSparseMatrix<int, T> m;
  CsrHold(int h, int w, int nnz, const int* rp, const int* ci, const T* v) {
    m.height = h; m.width = w; m.nnz = nnz;
    m.csrRowPtr = const_cast<int*>(rp);
    m.csrColIdx = const_cast<int*>(ci);
    m.csrVal = const_cast<T*>(v);
  }
  ~CsrHold() { m.csrRowPtr = NULL; m.csrColIdx = NULL; m.csrVal = NULL; }  // device copies freed by ~m
};
template <typename T>
struct VecHold {
  DenseVector<int, T> v;
  VecHold(int len, T* p) { v.length = len; v.val = p; v.val_gpu = NULL; v.n_gpu = 0; v.policy = none; }
  ~VecHold() { v.val = NULL; }
};
template <typename T>
struct MatHold {
  DenseMatrix<int, T> m;
  MatHold(int h, int w, T* p) { m.height = h; m.width = w; m.val = p; m.order = col_major; }
  ~MatHold() { m.val = NULL; }
};

double now_ms() {
  using namespace std::chrono;
  return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}
void sync_all(unsigned n_gpu) {
  for (unsigned i = 0; i < n_gpu; i++) { cudaSetDevice(i); cudaDeviceSynchronize(); }
  cudaSetDevice(0);
}

// y0 is re-uploaded into every replica before each repetition (outside the timed region), so
// every repetition computes the same y = beta*y0 + alpha*A*x; the last one is read back.
template <typename T>
int spmv_lane(int h, int w, int nnz, const int* rp, const int* ci, const T* v, const T* x, const T* y0,
              T* y_out, T alpha, T beta, unsigned n_gpu, int reps, double* ms_out) {
  CsrHold<T> A(h, w, nnz, rp, ci, v);
  VecHold<T> B(w, const_cast<T*>(x)), C(h, y_out);
  memcpy(y_out, y0, sizeof(T) * (size_t)h);
  A.m.sync2gpu(n_gpu, segment);      // spmv_test.cu:16-18
  B.v.sync2gpu(n_gpu, replicate);
  C.v.sync2gpu(n_gpu, replicate);
  for (int r = 0; r < reps; r++) {
    for (unsigned i = 0; i < n_gpu; i++) {
      cudaSetDevice(i);
      cudaMemcpy(C.v.val_gpu[i], y0, sizeof(T) * (size_t)h, cudaMemcpyHostToDevice);
    }
    sync_all(n_gpu);
    double t0 = now_ms();
    sblas_spmv_csr_v1<int, T>(&A.m, &B.v, &C.v, alpha, beta, n_gpu);   // spmv.h:36
    sync_all(n_gpu);
    ms_out[r] = now_ms() - t0;
  }
  C.v.sync2cpu(0);
  return 0;
}

template <typename T>
int spmm_lane(int method, int h, int w, int nnz, const int* rp, const int* ci, const T* v, int k,
              const T* Bv, const T* C0, T* C_out, T alpha, T beta, unsigned n_gpu, int reps, double* ms_out) {
  CsrHold<T> A(h, w, nnz, rp, ci, v);
  MatHold<T> B(w, k, const_cast<T*>(Bv)), C(h, k, C_out);
  size_t cbytes = sizeof(T) * (size_t)h * k;
  memcpy(C_out, C0, cbytes);
  if (method == 1) {                  // spmm_test.cu:28-30
    A.m.sync2gpu(n_gpu, replicate);
    B.m.sync2gpu(n_gpu, segment);
    C.m.sync2gpu(n_gpu, segment);
  } else {                            // spmm_test.cu:68-70
    A.m.sync2gpu(n_gpu, segment);
    B.m.sync2gpu(n_gpu, replicate);
    C.m.sync2gpu(n_gpu, replicate);
  }
  int avg = (int)ceil((double)k / n_gpu);   // matrix.h:560
  for (int r = 0; r < reps; r++) {
    for (unsigned i = 0; i < n_gpu; i++) {
      cudaSetDevice(i);
      if (method == 1)
        cudaMemcpy(C.m.val_gpu[i], C0 + (size_t)i * avg * h, sizeof(T) * (size_t)h * C.m.dim_gpu[i],
                   cudaMemcpyHostToDevice);
      else
        cudaMemcpy(C.m.val_gpu[i], C0, cbytes, cudaMemcpyHostToDevice);
    }
    sync_all(n_gpu);
    double t0 = now_ms();
    if (method == 1) sblas_spmm_csr_v1<int, T>(&A.m, &B.m, &C.m, alpha, beta, n_gpu);   // spmm.h:84
    else sblas_spmm_csr_v2<int, T>(&A.m, &B.m, &C.m, alpha, beta, n_gpu);               // spmm.h:164
    sync_all(n_gpu);
    ms_out[r] = now_ms() - t0;
  }
  if (method == 2) C.m.sync2cpu(0);   // v1 refreshes the host copy itself (spmm.h:151)
  return 0;
}
}  // namespace

extern "C" {
int refgpu_spmv_f64(int h, int w, int nnz, const int* rp, const int* ci, const double* v, const double* x,
                    const double* y0, double* y_out, double alpha, double beta, unsigned n_gpu, int reps,
                    double* ms_out) {
  return spmv_lane<double>(h, w, nnz, rp, ci, v, x, y0, y_out, alpha, beta, n_gpu, reps, ms_out);
}
int refgpu_spmv_f32(int h, int w, int nnz, const int* rp, const int* ci, const float* v, const float* x,
                    const float* y0, float* y_out, float alpha, float beta, unsigned n_gpu, int reps,
                    double* ms_out) {
  return spmv_lane<float>(h, w, nnz, rp, ci, v, x, y0, y_out, alpha, beta, n_gpu, reps, ms_out);
}
// fp64 only for method 2: the reference hard-codes ncclDouble there (spmm.h:261).
int refgpu_spmm_f64(int method, int h, int w, int nnz, const int* rp, const int* ci, const double* v, int k,
                    const double* B, const double* C0, double* C_out, double alpha, double beta,
                    unsigned n_gpu, int reps, double* ms_out) {
  return spmm_lane<double>(method, h, w, nnz, rp, ci, v, k, B, C0, C_out, alpha, beta, n_gpu, reps, ms_out);
}

// ---- library-kernel-only lane -----------------------------------------------------------------
// The same cuSPARSE calls the reference issues (descriptors and algorithm exactly as spmv.h:70-104
// and spmm.h:199-250: CSR int32 base 0, ALG_DEFAULT, column-major B / C with ld = height), but on
// operands that are ALREADY on the device, with the descriptor / buffer set-up outside the timed
// region and CUDA events on the handle's stream around the cusparseSpMV / cusparseSpMM call alone.
// This is the number a kernel-for-kernel comparison needs (ours: resident operands, CUDA events);
// the lanes above time what a caller of the reference's op sees (communicator + handle creation,
// zero-filled C_copy upload, all-reduce, axpby).  All pointers are DEVICE pointers.
#define LIBCHK(x) do { if ((x) != CUSPARSE_STATUS_SUCCESS) return -2; } while (0)
int libgpu_cusparse_spmv(int f64, long rows, long cols, long nnz, int* d_rp, int* d_ci, void* d_v, void* d_x,
                         void* d_y, double alpha, double beta, int warm, int reps, float* ms_out) {
  cudaDataType vt = f64 ? CUDA_R_64F : CUDA_R_32F;
  cusparseHandle_t h; LIBCHK(cusparseCreate(&h));
  cusparseSpMatDescr_t A; cusparseDnVecDescr_t X, Y;
  LIBCHK(cusparseCreateCsr(&A, rows, cols, nnz, d_rp, d_ci, d_v, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                           CUSPARSE_INDEX_BASE_ZERO, vt));
  LIBCHK(cusparseCreateDnVec(&X, cols, d_x, vt));
  LIBCHK(cusparseCreateDnVec(&Y, rows, d_y, vt));
  float af = (float)alpha, bf = (float)beta;
  const void* pa = f64 ? (const void*)&alpha : (const void*)&af;
  const void* pb = f64 ? (const void*)&beta : (const void*)&bf;
  size_t bs = 0; void* buf = NULL;
  LIBCHK(cusparseSpMV_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, pa, A, X, pb, Y, vt,
                                 CUSPARSE_SPMV_ALG_DEFAULT, &bs));
  if (cudaMalloc(&buf, bs ? bs : 16) != cudaSuccess) return -3;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int r = 0; r < warm + reps; r++) {
    if (r >= warm) cudaEventRecord(e0, 0);
    LIBCHK(cusparseSpMV(h, CUSPARSE_OPERATION_NON_TRANSPOSE, pa, A, X, pb, Y, vt, CUSPARSE_SPMV_ALG_DEFAULT, buf));
    if (r >= warm) { cudaEventRecord(e1, 0); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms_out[r - warm], e0, e1); }
  }
  cudaDeviceSynchronize();
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf);
  cusparseDestroySpMat(A); cusparseDestroyDnVec(X); cusparseDestroyDnVec(Y); cusparseDestroy(h);
  return cudaGetLastError() == cudaSuccess ? 0 : -4;
}
// B [cols x k] and C [rows x k] column-major (ldb = cols, ldc = rows), as the reference passes them.
int libgpu_cusparse_spmm(int f64, long rows, long cols, long nnz, int* d_rp, int* d_ci, void* d_v, long k,
                         void* d_B, void* d_C, double alpha, double beta, int warm, int reps, float* ms_out) {
  cudaDataType vt = f64 ? CUDA_R_64F : CUDA_R_32F;
  cusparseHandle_t h; LIBCHK(cusparseCreate(&h));
  cusparseSpMatDescr_t A; cusparseDnMatDescr_t B, C;
  LIBCHK(cusparseCreateCsr(&A, rows, cols, nnz, d_rp, d_ci, d_v, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                           CUSPARSE_INDEX_BASE_ZERO, vt));
  LIBCHK(cusparseCreateDnMat(&B, cols, k, cols, d_B, vt, CUSPARSE_ORDER_COL));
  LIBCHK(cusparseCreateDnMat(&C, rows, k, rows, d_C, vt, CUSPARSE_ORDER_COL));
  float af = (float)alpha, bf = (float)beta;
  const void* pa = f64 ? (const void*)&alpha : (const void*)&af;
  const void* pb = f64 ? (const void*)&beta : (const void*)&bf;
  size_t bs = 0; void* buf = NULL;
  LIBCHK(cusparseSpMM_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, pa, A, B,
                                 pb, C, vt, CUSPARSE_SPMM_ALG_DEFAULT, &bs));
  if (cudaMalloc(&buf, bs ? bs : 16) != cudaSuccess) return -3;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int r = 0; r < warm + reps; r++) {
    if (r >= warm) cudaEventRecord(e0, 0);
    LIBCHK(cusparseSpMM(h, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, pa, A, B, pb, C, vt,
                        CUSPARSE_SPMM_ALG_DEFAULT, buf));
    if (r >= warm) { cudaEventRecord(e1, 0); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms_out[r - warm], e0, e1); }
  }
  cudaDeviceSynchronize();
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf);
  cusparseDestroySpMat(A); cusparseDestroyDnMat(B); cusparseDestroyDnMat(C); cusparseDestroy(h);
  return cudaGetLastError() == cudaSuccess ? 0 : -4;
}
}  // extern "C"
