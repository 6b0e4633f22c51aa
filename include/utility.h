/* include/utility.h — helpers of the drop-in headers (B200 build).
 *
 * Keeps the names a user of the reference's utility.h relies on on the SpMV/SpMM path:
 *   CUDA_SAFE_CALL / CUDA_CHECK_ERROR            (utility.h:27-58)
 *   SAFE_ALOC_HOST / SAFE_ALOC_GPU / SAFE_FREE_* (utility.h:86-127)
 *   print_1d_array, cpu_timer, gpu_timer          (utility.h:130-178)
 *   check_equal, rand0to1                         (utility.h:182-197)
 *   exclusive_scan, csr_findRowIdxUsingNnzIdx     (utility.h:276-300)
 * but every allocation / copy goes through the C ABI of libsblas_b200.so, so this header
 * (and matrix.h, spmv.h, spmm.h on top of it) compiles with a plain C++ compiler; when the CUDA
 * runtime header is visible (nvcc, or g++ -I$CUDA/include) the CUDA_* macros additionally
 * accept cudaError_t expressions exactly like the reference's.
 * Not carried over: the key/value quicksort, segmented_sum / reduce_sum and sum_32_shfl
 * (utility.h:201-272: never called on this path) and the cuSPARSE/NCCL type mappers and
 * CHECK_CUSPARSE / CHECK_NCCL (utility.h:60-83, :302-316: there is no cuSPARSE or NCCL call
 * left on the path). */
#ifndef UTILITY_H
#define UTILITY_H

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "config.h"
#include "sblas_b200.h"

#if defined(__CUDACC__) || defined(SBLAS_WITH_CUDA_RUNTIME)
#define SBLAS_HAVE_CUDA_RUNTIME 1
#elif defined(__has_include)
#if __has_include(<cuda_runtime.h>)
#define SBLAS_HAVE_CUDA_RUNTIME 1
#endif
#endif
#ifdef SBLAS_HAVE_CUDA_RUNTIME
#include <cuda_runtime.h>
#endif

/* ---- error handling: the C ABI returns codes, the reference prints and exits ------------- */
#define SBLAS_SAFE_CALL(rc) sblas_safe_call_((rc), __FILE__, __LINE__)
inline void sblas_safe_call_(int rc, const char* file, int line) {
#ifdef CUDA_ERROR_CHECK
  if (rc != 0) {
    fprintf(stderr, "sblas call failed at %s:%i : %s\n", file, line, sblas_last_error());
    exit(-1);
  }
#else
  (void)rc; (void)file; (void)line;
#endif
}

#ifdef SBLAS_HAVE_CUDA_RUNTIME
#define CUDA_SAFE_CALL(err) sblas_cuda_safe_call_((err), __FILE__, __LINE__)
inline void sblas_cuda_safe_call_(cudaError_t err, const char* file, int line) {
#ifdef CUDA_ERROR_CHECK
  if (err != cudaSuccess) {
    fprintf(stderr, "cudaSafeCall() failed at %s:%i : %s\n", file, line, cudaGetErrorString(err));
    exit(-1);
  }
#endif
}
#define CUDA_CHECK_ERROR() sblas_cuda_check_error_(__FILE__, __LINE__)
inline void sblas_cuda_check_error_(const char* file, int line) {
#ifdef CUDA_ERROR_CHECK
  cudaError_t err = cudaGetLastError();
  if (err == cudaSuccess) err = cudaDeviceSynchronize();
  if (err != cudaSuccess) {
    fprintf(stderr, "cudaCheckError() failed at %s:%i : %s\n", file, line, cudaGetErrorString(err));
    exit(-1);
  }
#endif
}
#else
/* no CUDA runtime header: only the "sync and check" use of the macros is meaningful */
#define CUDA_SAFE_CALL(rc) SBLAS_SAFE_CALL(rc)
#define CUDA_CHECK_ERROR() SBLAS_SAFE_CALL(sblas_dev_sync(0))
#endif

/* synchronise logical GPUs 0..n-1 (what `CUDA_SAFE_CALL(cudaDeviceSynchronize())` does for the
 * current device in the reference's drivers, spmv_test.cu:19) */
inline void sblas_sync_gpus(unsigned n_gpu) {
  for (unsigned i = 0; i < n_gpu; ++i) SBLAS_SAFE_CALL(sblas_dev_sync((int)i));
}

/* ---- allocation (same macro names; pinned host memory / device memory of logical GPU i) --- */
#define SAFE_ALOC_HOST(X, Y) SBLAS_SAFE_CALL(sblas_host_alloc((size_t)(Y), (void**)&(X)))
#define SAFE_FREE_HOST(X)                       \
  if ((X) != NULL) {                            \
    SBLAS_SAFE_CALL(sblas_host_free((void*)(X))); \
    (X) = NULL;                                 \
  }
/* the reference's SAFE_ALOC_GPU allocates on the *current* device after cudaSetDevice(i);
 * here the logical GPU is an explicit argument */
#define SAFE_ALOC_GPU_ON(I, X, Y) SBLAS_SAFE_CALL(sblas_dev_alloc((int)(I), (size_t)(Y), (void**)&(X)))
#define SAFE_FREE_GPU_ON(I, X)                              \
  if ((X) != NULL) {                                        \
    SBLAS_SAFE_CALL(sblas_dev_free((int)(I), (void*)(X)));  \
    (X) = NULL;                                             \
  }
#define SAFE_FREE_MULTI_GPU(X, Y)                                               \
  if ((X) != NULL) {                                                            \
    for (unsigned i__ = 0; i__ < (Y); i__++)                                    \
      if ((X)[i__] != NULL) SBLAS_SAFE_CALL(sblas_dev_free((int)i__, (void*)(X)[i__])); \
    SBLAS_SAFE_CALL(sblas_host_free((void*)(X)));                               \
    (X) = NULL;                                                                 \
  }
#define SAFE_ALOC_MANAGED(X, Y) SBLAS_SAFE_CALL(sblas_managed_alloc((size_t)(Y), (void**)&(X)))
#define SAFE_FREE_MANAGED(X)                           \
  if ((X) != NULL) {                                   \
    SBLAS_SAFE_CALL(sblas_managed_free((void*)(X)));   \
    (X) = NULL;                                        \
  }

/* ---- printing ------------------------------------------------------------------------------ */
template <typename T>
void print_1d_array(const T* input, int length) {
  for (int i = 0; i < length; i++) printf("%.3lf, %s", (double)input[i], ((i + 1) % 10 == 0) ? "\n" : "");
  printf("\n");
}

/* ---- timers (milliseconds) ----------------------------------------------------------------- */
inline double get_cpu_timer() {
  using clk = std::chrono::system_clock;
  return std::chrono::duration<double, std::milli>(clk::now().time_since_epoch()).count();
}
typedef struct CPU_Timer {
  double start = 0.0, stop = 0.0;
  void start_timer() { start = get_cpu_timer(); }
  void stop_timer() { stop = get_cpu_timer(); }
  double measure() const { return stop - start; }
} cpu_timer;

#ifdef SBLAS_HAVE_CUDA_RUNTIME
typedef struct GPU_Timer { /* CUDA events on the current device's default stream (utility.h:162-178) */
  cudaEvent_t start, stop;
  GPU_Timer() {
    CUDA_SAFE_CALL(cudaEventCreate(&start));
    CUDA_SAFE_CALL(cudaEventCreate(&stop));
  }
  ~GPU_Timer() {
    cudaEventDestroy(start);
    cudaEventDestroy(stop);
  }
  void start_timer() { CUDA_SAFE_CALL(cudaEventRecord(start)); }
  void stop_timer() { CUDA_SAFE_CALL(cudaEventRecord(stop)); }
  double measure() {
    float ms = 0.f;
    CUDA_SAFE_CALL(cudaEventSynchronize(stop));
    CUDA_SAFE_CALL(cudaEventElapsedTime(&ms, start, stop));
    return (double)ms;
  }
} gpu_timer;
#else
typedef struct GPU_Timer { /* device-synchronised wall clock when no CUDA header is available */
  double start = 0.0, stop = 0.0;
  void start_timer() { SBLAS_SAFE_CALL(sblas_dev_sync(0)); start = get_cpu_timer(); }
  void stop_timer() { SBLAS_SAFE_CALL(sblas_dev_sync(0)); stop = get_cpu_timer(); }
  double measure() const { return stop - start; }
} gpu_timer;
#endif

/* ---- verification (utility.h:182-193): absolute difference against ERROR_BAR --------------- */
template <typename T>
bool check_equal(const T* x, const T* y, size_t m) {
  size_t bad = 0;
  for (size_t i = 0; i < m; i++) bad += (std::fabs((double)x[i] - (double)y[i]) > ERROR_BAR) ? 1 : 0;
  return bad == 0;
}
/* The stricter per-row bound of the B200 build's own tests: |x-y| <= rel * bound[i]. */
template <typename T>
bool check_equal_scaled(const T* x, const T* y, const T* bound, size_t m, double rel) {
  for (size_t i = 0; i < m; i++)
    if (std::fabs((double)x[i] - (double)y[i]) > rel * (double)bound[i]) return false;
  return true;
}

/* utility.h:197 — glibc rand(), seeded by the constructors with RAND_INIT_SEED */
inline double rand0to1() { return (double)rand() / (double)RAND_MAX; }

/* utility.h:276-288 — in-place exclusive prefix sum */
template <typename IdxType, typename DataType>
void exclusive_scan(DataType* input, IdxType length) {
  DataType running = 0;
  for (IdxType i = 0; i < length; i++) {
    const DataType v = input[i];
    input[i] = running;
    running += v;
  }
}

/* utility.h:293-300 — the (non-empty) row that holds nnz index nnzIdx, -1 if none.  The
 * reference scans all rows; this is upper_bound - 1 on the monotone row-pointer array. */
template <typename IdxType>
IdxType csr_findRowIdxUsingNnzIdx(const IdxType* rowPtr, IdxType height, IdxType nnzIdx) {
  if (sizeof(IdxType) == 4 && height > 0 && rowPtr[0] >= 0)
    return (IdxType)sblas_csr_find_row_i32((const int32_t*)rowPtr, (int32_t)height, (int32_t)nnzIdx);
  IdxType lo = 0, hi = height + 1; /* first index with rowPtr[idx] > nnzIdx */
  while (lo < hi) {
    const IdxType mid = lo + (hi - lo) / 2;
    if (rowPtr[mid] > nnzIdx) hi = mid; else lo = mid + 1;
  }
  if (lo == 0 || lo > height) return (IdxType)-1;
  return lo - 1;
}

/* ---- f32 / f64 dispatch onto the C ABI ------------------------------------------------------ */
/* scoped NVTX range (sblas_range_push / _pop) */
struct sblas_range {
  explicit sblas_range(const char* name) { sblas_range_push(name); }
  ~sblas_range() { sblas_range_pop(); }
  sblas_range(const sblas_range&) = delete;
  sblas_range& operator=(const sblas_range&) = delete;
};

template <typename DataType> struct sblas_abi; /* only float and double (utility.h:304-306) */
template <> struct sblas_abi<double> {
  static int upload(int g, const double* h, int64_t second, int64_t dim, int32_t lay, double** d, int64_t* ld) { return sblas_dense_upload_f64(g, h, second, dim, lay, d, ld); }
  static int download(int g, const double* d, int64_t ld, int64_t second, int64_t dim, int32_t lay, double* h) { return sblas_dense_download_f64(g, d, ld, second, dim, lay, h); }
  static int download_multi(unsigned n, const double* const* d, const int64_t* ld, int64_t second, const int64_t* dim, int32_t lay, double* const* h) { return sblas_dense_download_multi_f64(n, d, ld, second, dim, lay, h); }
  static int axpby(double* v0, const double* v1, double a, double b, int64_t n) { return sblas_axpby_f64(NULL, v0, v1, a, b, n); }
  static int spmv_v1(unsigned n, int32_t h, int32_t w, const int32_t* const* rp, const int32_t* const* ci, const double* const* v, const int32_t* ng, const int32_t* s, const int32_t* e, const double* const* x, double* const* y, double a, double b, const sblasPlan_t* p) { return sblas_mgpu_spmv_csr_v1_f64(n, h, w, rp, ci, v, ng, s, e, x, y, a, b, p); }
  static int spmm_v1(unsigned n, int32_t h, int32_t w, int32_t nnz, const int32_t* const* rp, const int32_t* const* ci, const double* const* v, const int32_t* dim, int32_t lay, const double* const* B, const int64_t* ldb, double* const* C, const int64_t* ldc, double a, double b, const sblasPlan_t* p) { return sblas_mgpu_spmm_csr_v1_f64(n, h, w, nnz, rp, ci, v, dim, lay, B, ldb, C, ldc, a, b, p); }
  static int spmm_v2(unsigned n, int32_t h, int32_t w, const int32_t* const* rp, const int32_t* const* ci, const double* const* v, const int32_t* ng, const int32_t* s, const int32_t* e, int32_t k, int32_t lay, const double* const* B, int64_t ldb, double* const* C, int64_t ldc, double a, double b, const sblasPlan_t* p) { return sblas_mgpu_spmm_csr_v2_f64(n, h, w, rp, ci, v, ng, s, e, k, lay, B, ldb, C, ldc, a, b, p); }
};
template <> struct sblas_abi<float> {
  static int upload(int g, const float* h, int64_t second, int64_t dim, int32_t lay, float** d, int64_t* ld) { return sblas_dense_upload_f32(g, h, second, dim, lay, d, ld); }
  static int download(int g, const float* d, int64_t ld, int64_t second, int64_t dim, int32_t lay, float* h) { return sblas_dense_download_f32(g, d, ld, second, dim, lay, h); }
  static int download_multi(unsigned n, const float* const* d, const int64_t* ld, int64_t second, const int64_t* dim, int32_t lay, float* const* h) { return sblas_dense_download_multi_f32(n, d, ld, second, dim, lay, h); }
  static int axpby(float* v0, const float* v1, float a, float b, int64_t n) { return sblas_axpby_f32(NULL, v0, v1, a, b, n); }
  static int spmv_v1(unsigned n, int32_t h, int32_t w, const int32_t* const* rp, const int32_t* const* ci, const float* const* v, const int32_t* ng, const int32_t* s, const int32_t* e, const float* const* x, float* const* y, float a, float b, const sblasPlan_t* p) { return sblas_mgpu_spmv_csr_v1_f32(n, h, w, rp, ci, v, ng, s, e, x, y, a, b, p); }
  static int spmm_v1(unsigned n, int32_t h, int32_t w, int32_t nnz, const int32_t* const* rp, const int32_t* const* ci, const float* const* v, const int32_t* dim, int32_t lay, const float* const* B, const int64_t* ldb, float* const* C, const int64_t* ldc, float a, float b, const sblasPlan_t* p) { return sblas_mgpu_spmm_csr_v1_f32(n, h, w, nnz, rp, ci, v, dim, lay, B, ldb, C, ldc, a, b, p); }
  static int spmm_v2(unsigned n, int32_t h, int32_t w, const int32_t* const* rp, const int32_t* const* ci, const float* const* v, const int32_t* ng, const int32_t* s, const int32_t* e, int32_t k, int32_t lay, const float* const* B, int64_t ldb, float* const* C, int64_t ldc, float a, float b, const sblasPlan_t* p) { return sblas_mgpu_spmm_csr_v2_f32(n, h, w, rp, ci, v, ng, s, e, k, lay, B, ldb, C, ldc, a, b, p); }
};

#endif /* UTILITY_H */
