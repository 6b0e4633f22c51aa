// memory.cu — memory plumbing behind the C++ drop-in headers (include/matrix.h, matrix_um.h).
//
// The reference's data model allocates pinned host arrays and per-GPU device arrays with the
// macros of utility.h:86-127 and moves them with plain cudaMemcpy (matrix.h:341-387, :541-570,
// :595-612, :688-713); the managed-memory variant uses cudaMallocManaged + cudaMemAdvise +
// cudaMemPrefetchAsync (matrix_um.h:529-614).  The drop-in headers make the same moves
// through these entry points, so they compile with a plain C++ compiler (no CUDA toolkit on
// the include path) and every CUDA call of the product lives inside libsblas_b200.so.
// `gpu` is always a LOGICAL GPU ordinal (sblas_device_of).
#include <nvtx3/nvToolsExt.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace sblas {

static int g_has_device = -1;  // decided once: pinned host memory needs a driver
static bool has_device() {
  if (g_has_device < 0) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
    g_has_device = n > 0 ? 1 : 0;
  }
  return g_has_device == 1;
}

struct DeviceScope {  // cudaSetDevice(device_of(gpu)) for the lifetime of the object
  int prev = -1;
  bool ok = true;
  explicit DeviceScope(int gpu) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (cudaSetDevice(sblas_device_of(gpu)) != cudaSuccess) ok = false;
  }
  ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};

#define SBLAS_ON_GPU(gpu)                                                                  \
  DeviceScope scope__(gpu);                                                                \
  if (!scope__.ok) {                                                                       \
    set_error("cannot select logical GPU %d (device %d): %s", gpu, sblas_device_of(gpu),   \
              cudaGetErrorString(cudaGetLastError()));                                     \
    return 1;                                                                              \
  }

template <typename T>
static int dense_upload(int gpu, const T* h_src, int64_t second, int64_t dim, int32_t layout,
                        T** d_out, int64_t* ld_out) {
  SBLAS_CHECK(d_out && ld_out && second >= 0 && dim >= 0, "dense_upload: bad arguments");
  *d_out = nullptr;
  *ld_out = layout == SBLAS_ROW_MAJOR ? 1 : second;
  if (second == 0 || dim == 0) return 0;
  SBLAS_CHECK(h_src, "dense_upload: null source");
  SBLAS_ON_GPU(gpu);
  const size_t bytes = sizeof(T) * (size_t)second * (size_t)dim;
  T* stage = nullptr;
  SBLAS_CUDA(cudaMalloc(&stage, bytes));
  SBLAS_CUDA(cudaMemcpy(stage, h_src, bytes, cudaMemcpyHostToDevice));
  if (layout == SBLAS_COL_MAJOR) {  // the reference's device layout: slice as is
    *d_out = stage;
    *ld_out = second;
    return 0;
  }
  // internal layout: row-major [second x ld], ld = dim rounded up to a 16-byte multiple
  constexpr int64_t VF = 16 / sizeof(T);
  const int64_t ld = (dim + VF - 1) / VF * VF;
  T* dst = nullptr;
  SBLAS_CUDA(cudaMalloc(&dst, sizeof(T) * (size_t)second * (size_t)ld));
  if (ld != dim) SBLAS_CUDA(cudaMemset(dst, 0, sizeof(T) * (size_t)second * (size_t)ld));
  int rc = sizeof(T) == 8
               ? sblas_transpose_f64(nullptr, (const double*)stage, second, dim, second, (double*)dst, ld)
               : sblas_transpose_f32(nullptr, (const float*)stage, second, dim, second, (float*)dst, ld);
  if (rc) return rc;
  SBLAS_CUDA(cudaDeviceSynchronize());
  SBLAS_CUDA(cudaFree(stage));
  *d_out = dst;
  *ld_out = ld;
  return 0;
}

template <typename T>
static int dense_download(int gpu, const T* d_src, int64_t ld, int64_t second, int64_t dim,
                          int32_t layout, T* h_dst) {
  if (second <= 0 || dim <= 0) return 0;
  SBLAS_CHECK(d_src && h_dst, "dense_download: null pointer");
  SBLAS_ON_GPU(gpu);
  const size_t bytes = sizeof(T) * (size_t)second * (size_t)dim;
  if (layout == SBLAS_COL_MAJOR) {
    SBLAS_CHECK(ld == second, "dense_download: col-major slices are contiguous (ld == rows)");
    SBLAS_CUDA(cudaMemcpy(h_dst, d_src, bytes, cudaMemcpyDeviceToHost));
    return 0;
  }
  T* stage = nullptr;
  SBLAS_CUDA(cudaMalloc(&stage, bytes));
  int rc = sizeof(T) == 8
               ? sblas_transpose_f64(nullptr, (const double*)d_src, ld, second, dim, (double*)stage, second)
               : sblas_transpose_f32(nullptr, (const float*)d_src, ld, second, dim, (float*)stage, second);
  if (rc) return rc;
  SBLAS_CUDA(cudaMemcpy(h_dst, stage, bytes, cudaMemcpyDeviceToHost));
  SBLAS_CUDA(cudaFree(stage));
  return 0;
}

// ---- streamed sync2gpu / sync2cpu of a row range (caller-owned device storage) -----------------------
// DenseMatrix::sync2gpu / sync2cpu move a col-major host operand (matrix.h:541-570, :595-612); the device
// keeps it row-major (see dense_upload above).  These variants are what a caller that re-sends B / C on
// every step uses: no allocation, no synchronisation, and the PCIe copy of chunk i+1 overlaps the layout
// conversion of chunk i.  A chunk is a range of rows: k pieces of (rows * sizeof(T)) contiguous bytes on
// the host, one strided DMA (cudaMemcpy2DAsync) into a staging buffer [k x rows], then the transpose kernel
// into the row-major destination.  Copies run on a per-device helper stream (one per direction, so H2D and
// D2H overlap on the full-duplex link), conversions on the caller's stream; events order the two.
struct XferCtx {
  cudaStream_t copy = nullptr;
  void* stage[2] = {nullptr, nullptr};
  cudaEvent_t ready[2] = {nullptr, nullptr}, free_[2] = {nullptr, nullptr}, done = nullptr;
  unsigned turn = 0;
};
constexpr size_t kXferStageBytes = 128u << 20;
static XferCtx g_up[SBLAS_MAX_GPU * 4], g_down[SBLAS_MAX_GPU * 4];  // indexed by CUDA device

static int xfer_get(XferCtx* table, XferCtx** out) {
  int dev = 0;
  SBLAS_CUDA(cudaGetDevice(&dev));
  SBLAS_CHECK(dev >= 0 && dev < SBLAS_MAX_GPU * 4, "device ordinal %d out of range", dev);
  XferCtx& c = table[dev];
  if (!c.copy) {
    SBLAS_CUDA(cudaStreamCreateWithFlags(&c.copy, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      SBLAS_CUDA(cudaMalloc(&c.stage[b], kXferStageBytes));
      SBLAS_CUDA(cudaEventCreateWithFlags(&c.ready[b], cudaEventDisableTiming));
      SBLAS_CUDA(cudaEventCreateWithFlags(&c.free_[b], cudaEventDisableTiming));
    }
    SBLAS_CUDA(cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming));
  }
  *out = &c;
  return 0;
}

template <typename T>
static int transpose_on(cudaStream_t st, const T* in, int64_t in_ld, int64_t n_major, int64_t n_minor, T* out, int64_t out_ld) {
  return sizeof(T) == 8 ? sblas_transpose_f64(st, (const double*)in, in_ld, n_major, n_minor, (double*)out, out_ld)
                        : sblas_transpose_f32(st, (const float*)in, in_ld, n_major, n_minor, (float*)out, out_ld);
}

template <typename T>
static int upload_rows(cudaStream_t st, const T* h, int64_t h_ld, int64_t r0, int64_t nr, int32_t k, T* d_dst, int64_t ld) {
  SBLAS_CHECK(nr >= 0 && k >= 0 && h_ld >= 0 && r0 >= 0 && ld >= k, "dense_upload_rows: bad arguments");
  if (nr == 0 || k == 0) return 0;
  SBLAS_CHECK(h && d_dst, "dense_upload_rows: null pointer");
  XferCtx* c = nullptr;
  if (xfer_get(g_up, &c)) return 1;
  const int64_t chunk = std::max<int64_t>(1, (int64_t)(kXferStageBytes / sizeof(T)) / k);
  for (int64_t r = 0; r < nr; r += chunk) {
    const int64_t cr = std::min(chunk, nr - r);
    const int b = (int)(c->turn++ & 1u);
    SBLAS_CUDA(cudaStreamWaitEvent(c->copy, c->free_[b], 0));   // the conversion that last read this buffer
    SBLAS_CUDA(cudaMemcpy2DAsync(c->stage[b], (size_t)cr * sizeof(T), h + r0 + r, (size_t)h_ld * sizeof(T),
                                 (size_t)cr * sizeof(T), (size_t)k, cudaMemcpyHostToDevice, c->copy));
    SBLAS_CUDA(cudaEventRecord(c->ready[b], c->copy));
    SBLAS_CUDA(cudaStreamWaitEvent(st, c->ready[b], 0));
    if (int rc = transpose_on<T>(st, (const T*)c->stage[b], cr, k, cr, d_dst + r * ld, ld)) return rc;
    SBLAS_CUDA(cudaEventRecord(c->free_[b], st));
  }
  return 0;
}

template <typename T>
static int download_rows(cudaStream_t st, const T* d_src, int64_t ld, int64_t nr, int32_t k, T* h, int64_t h_ld, int64_t r0) {
  SBLAS_CHECK(nr >= 0 && k >= 0 && h_ld >= 0 && r0 >= 0 && ld >= k, "dense_download_rows: bad arguments");
  if (nr == 0 || k == 0) return 0;
  SBLAS_CHECK(h && d_src, "dense_download_rows: null pointer");
  XferCtx* c = nullptr;
  if (xfer_get(g_down, &c)) return 1;
  const int64_t chunk = std::max<int64_t>(1, (int64_t)(kXferStageBytes / sizeof(T)) / k);
  for (int64_t r = 0; r < nr; r += chunk) {
    const int64_t cr = std::min(chunk, nr - r);
    const int b = (int)(c->turn++ & 1u);
    SBLAS_CUDA(cudaStreamWaitEvent(st, c->free_[b], 0));        // the copy that last drained this buffer
    if (int rc = transpose_on<T>(st, d_src + r * ld, ld, cr, k, (T*)c->stage[b], cr)) return rc;
    SBLAS_CUDA(cudaEventRecord(c->ready[b], st));
    SBLAS_CUDA(cudaStreamWaitEvent(c->copy, c->ready[b], 0));
    SBLAS_CUDA(cudaMemcpy2DAsync(h + r0 + r, (size_t)h_ld * sizeof(T), c->stage[b], (size_t)cr * sizeof(T),
                                 (size_t)cr * sizeof(T), (size_t)k, cudaMemcpyDeviceToHost, c->copy));
    SBLAS_CUDA(cudaEventRecord(c->free_[b], c->copy));
  }
  // the caller's stream completes only when the last bytes are on the host
  SBLAS_CUDA(cudaEventRecord(c->done, c->copy));
  SBLAS_CUDA(cudaStreamWaitEvent(st, c->done, 0));
  return 0;
}

// DenseMatrix::sync2cpu of EVERY GPU's slice at once (what sblas_spmm_csr_v1 does after its kernels, spmm.h:151):
// the downloads of all GPUs are enqueued first (each on its own device: layout conversion chunks + D2H copies over
// that GPU's own host link), then all devices are synchronised — n host links in parallel instead of n serial
// cudaMemcpy calls with a staging cudaMalloc / cudaFree each.
template <typename T>
static int dense_download_multi(unsigned n_gpu, const T* const* d, const int64_t* ld, int64_t second, const int64_t* dim,
                                int32_t layout, T* const* h) {
  SBLAS_CHECK(n_gpu >= 1 && n_gpu <= SBLAS_MAX_GPU && d && ld && dim && h, "dense_download_multi: bad arguments");
  int rc = 0;
  for (unsigned i = 0; i < n_gpu && rc == 0; ++i) {
    if (dim[i] <= 0 || second <= 0) continue;
    SBLAS_CHECK(d[i] && h[i], "dense_download_multi: null pointer for GPU %u", i);
    SBLAS_ON_GPU((int)i);
    if (layout == SBLAS_COL_MAJOR) {
      SBLAS_CHECK(ld[i] == second, "dense_download_multi: col-major slices are contiguous (ld == rows)");
      SBLAS_CUDA(cudaMemcpyAsync(h[i], d[i], sizeof(T) * (size_t)second * (size_t)dim[i], cudaMemcpyDeviceToHost, nullptr));
    } else {
      rc = download_rows<T>(nullptr, d[i], ld[i], second, (int32_t)dim[i], h[i], second, 0);
    }
  }
  for (unsigned i = 0; i < n_gpu; ++i) {
    if (dim[i] <= 0 || second <= 0) continue;
    SBLAS_ON_GPU((int)i);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess && rc == 0) { set_error("dense_download_multi: GPU %u: %s", i, cudaGetErrorString(e)); rc = 1; }
  }
  return rc;
}

}  // namespace sblas

using namespace sblas;
extern "C" {

int sblas_dense_download_multi_f64(unsigned n_gpu, const double* const* d, const int64_t* ld, int64_t second,
                                   const int64_t* dim, int32_t dev_layout, double* const* h) {
  return dense_download_multi<double>(n_gpu, d, ld, second, dim, dev_layout, h);
}
int sblas_dense_download_multi_f32(unsigned n_gpu, const float* const* d, const int64_t* ld, int64_t second,
                                   const int64_t* dim, int32_t dev_layout, float* const* h) {
  return dense_download_multi<float>(n_gpu, d, ld, second, dim, dev_layout, h);
}

int sblas_dense_upload_rows_f64(void* stream, const double* h, int64_t h_ld, int64_t r0, int64_t nr, int32_t k,
                                double* d_dst, int64_t ld) {
  return upload_rows<double>((cudaStream_t)stream, h, h_ld, r0, nr, k, d_dst, ld);
}
int sblas_dense_upload_rows_f32(void* stream, const float* h, int64_t h_ld, int64_t r0, int64_t nr, int32_t k,
                                float* d_dst, int64_t ld) {
  return upload_rows<float>((cudaStream_t)stream, h, h_ld, r0, nr, k, d_dst, ld);
}
int sblas_dense_download_rows_f64(void* stream, const double* d_src, int64_t ld, int64_t nr, int32_t k, double* h,
                                  int64_t h_ld, int64_t r0) {
  return download_rows<double>((cudaStream_t)stream, d_src, ld, nr, k, h, h_ld, r0);
}
int sblas_dense_download_rows_f32(void* stream, const float* d_src, int64_t ld, int64_t nr, int32_t k, float* h,
                                  int64_t h_ld, int64_t r0) {
  return download_rows<float>((cudaStream_t)stream, d_src, ld, nr, k, h, h_ld, r0);
}

// NVTX ranges (matrix_um.h:30, :435-445, :485-519 of the reference mark its load / distribute / tuning phases):
// nvtx3 is header-only and costs a null-pointer test per call unless a tool (nsys, ncu --nvtx) is attached
int sblas_range_push(const char* name) { return nvtxRangePushA(name ? name : "sblas"); }
int sblas_range_pop(void) { return nvtxRangePop(); }

int sblas_host_alloc(size_t bytes, void** p) {
  SBLAS_CHECK(p, "host_alloc: null out pointer");
  *p = nullptr;
  if (bytes == 0) bytes = 1;
  if (has_device()) {
    SBLAS_CUDA(cudaMallocHost(p, bytes));  // SAFE_ALOC_HOST, utility.h:122
  } else {
    // no driver (CPU-only host logic / partition tests): ordinary memory; no kernel can run anyway
    SBLAS_CHECK(posix_memalign(p, 64, bytes) == 0, "host_alloc: out of memory (%zu bytes)", bytes);
  }
  return 0;
}
int sblas_host_free(void* p) {
  if (!p) return 0;
  if (has_device()) SBLAS_CUDA(cudaFreeHost(p)); else free(p);
  return 0;
}
int sblas_dev_alloc(int gpu, size_t bytes, void** d_p) {
  SBLAS_CHECK(d_p, "dev_alloc: null out pointer");
  *d_p = nullptr;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaMalloc(d_p, bytes ? bytes : 16));  // SAFE_ALOC_GPU, utility.h:86
  return 0;
}
int sblas_dev_free(int gpu, void* d_p) {
  if (!d_p) return 0;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaFree(d_p));
  return 0;
}
int sblas_copy_h2d(int gpu, void* d_dst, const void* h_src, size_t bytes) {
  if (bytes == 0) return 0;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaMemcpy(d_dst, h_src, bytes, cudaMemcpyHostToDevice));
  return 0;
}
int sblas_copy_d2h(int gpu, void* h_dst, const void* d_src, size_t bytes) {
  if (bytes == 0) return 0;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaMemcpy(h_dst, d_src, bytes, cudaMemcpyDeviceToHost));
  return 0;
}
int sblas_copy_d2d(int gpu, void* d_dst, const void* d_src, size_t bytes) {
  if (bytes == 0) return 0;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaMemcpy(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice));
  return 0;
}
int sblas_dev_memset(int gpu, void* d_p, int value, size_t bytes) {
  if (bytes == 0) return 0;
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaMemset(d_p, value, bytes));
  return 0;
}
int sblas_dev_sync(int gpu) {
  SBLAS_ON_GPU(gpu);
  SBLAS_CUDA(cudaDeviceSynchronize());
  return 0;
}

int sblas_dense_upload_f64(int gpu, const double* h, int64_t second, int64_t dim, int32_t layout,
                           double** d_out, int64_t* ld_out) {
  return dense_upload<double>(gpu, h, second, dim, layout, d_out, ld_out);
}
int sblas_dense_upload_f32(int gpu, const float* h, int64_t second, int64_t dim, int32_t layout,
                           float** d_out, int64_t* ld_out) {
  return dense_upload<float>(gpu, h, second, dim, layout, d_out, ld_out);
}
int sblas_dense_download_f64(int gpu, const double* d, int64_t ld, int64_t second, int64_t dim,
                             int32_t layout, double* h) {
  return dense_download<double>(gpu, d, ld, second, dim, layout, h);
}
int sblas_dense_download_f32(int gpu, const float* d, int64_t ld, int64_t second, int64_t dim,
                             int32_t layout, float* h) {
  return dense_download<float>(gpu, d, ld, second, dim, layout, h);
}

// sblas_plan_create_i32 / sblas_axpby_* on a logical GPU (the classes of matrix.h never call
// cudaSetDevice themselves)
int sblas_plan_create_on(int gpu, int32_t rows, int64_t nnz, const int32_t* d_rowptr, sblasPlan_t* plan) {
  SBLAS_ON_GPU(gpu);
  return sblas_plan_create_i32(nullptr, rows, nnz, d_rowptr, plan);
}
int sblas_axpby_on(int gpu, int is_f64, void* d_v0, const void* d_v1, double alpha, double beta, int64_t n) {
  SBLAS_ON_GPU(gpu);
  return is_f64 ? sblas_axpby_f64(nullptr, (double*)d_v0, (const double*)d_v1, alpha, beta, n)
                : sblas_axpby_f32(nullptr, (float*)d_v0, (const float*)d_v1, (float)alpha, (float)beta, n);
}

int sblas_plan_build_rowblocks_on(int gpu, int is_f64, sblasPlan_t plan, const int32_t* d_colidx, const void* d_val) {
  if (!plan) return 0;
  SBLAS_ON_GPU(gpu);
  return is_f64 ? sblas_plan_build_rowblocks_f64(nullptr, plan, d_colidx, (const double*)d_val, 0)
                : sblas_plan_build_rowblocks_f32(nullptr, plan, d_colidx, (const float*)d_val, 0);
}
int sblas_plan_prepare_spmv_on(int gpu, int is_f64, sblasPlan_t plan, const int32_t* d_colidx, const void* d_val, int32_t cols) {
  if (!plan) return 0;
  SBLAS_ON_GPU(gpu);
  int rc = sblas_plan_prepare_i32(nullptr, plan, d_colidx, cols, is_f64 ? 8 : 4);
  if (rc) return rc;
  return is_f64 ? sblas_plan_build_colpanels_f64(nullptr, plan, d_colidx, (const double*)d_val, cols, 0)
                : sblas_plan_build_colpanels_f32(nullptr, plan, d_colidx, (const float*)d_val, cols, 0);
}
// CsrToCsc / CscToCsr / sortByRow (matrix.h:40-91, :221-237) on a logical GPU's default stream; for
// transpose_on, (m, n, ptr, idx) are (rows, cols, rowptr, colidx) for CSR -> CSC and (cols, rows,
// colptr, rowidx) for CSC -> CSR.
int sblas_compressed_transpose_on(int gpu, int is_f64, int32_t m, int32_t n, int64_t nnz, const int32_t* d_ptr,
                                  const int32_t* d_idx, const void* d_val, int32_t* d_out_ptr, int32_t* d_out_idx,
                                  void* d_out_val) {
  SBLAS_ON_GPU(gpu);
  return is_f64 ? sblas_csr2csc_f64(nullptr, m, n, nnz, d_ptr, d_idx, (const double*)d_val, d_out_ptr, d_out_idx, (double*)d_out_val)
                : sblas_csr2csc_f32(nullptr, m, n, nnz, d_ptr, d_idx, (const float*)d_val, d_out_ptr, d_out_idx, (float*)d_out_val);
}
int sblas_coo_sort_on(int gpu, int is_f64, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col, void* d_val) {
  SBLAS_ON_GPU(gpu);
  return is_f64 ? sblas_coo_sort_f64(nullptr, nnz, height, d_row, d_col, (double*)d_val)
                : sblas_coo_sort_f32(nullptr, nnz, height, d_row, d_col, (float*)d_val);
}
int sblas_scal_add_on(int gpu, int is_f64, void* d_v, double val, double beta, int64_t n) {
  SBLAS_ON_GPU(gpu);
  return is_f64 ? sblas_scal_add_f64(nullptr, (double*)d_v, val, beta, n)
                : sblas_scal_add_f32(nullptr, (float*)d_v, (float)val, (float)beta, n);
}

// ---- managed memory (matrix_um.h) ---------------------------------------------------------------
int sblas_managed_alloc(size_t bytes, void** p) {
  SBLAS_CHECK(p, "managed_alloc: null out pointer");
  *p = nullptr;
  SBLAS_CHECK(has_device(), "managed_alloc: no CUDA device");
  SBLAS_CUDA(cudaMallocManaged(p, bytes ? bytes : 16, cudaMemAttachGlobal));
  return 0;
}
int sblas_managed_free(void* p) {
  if (p) SBLAS_CUDA(cudaFree(p));
  return 0;
}
int sblas_mem_advise(void* p, size_t bytes, int advice, int gpu) {
  if (!p || bytes == 0) return 0;
  cudaMemoryAdvise a;
  switch (advice) {
    case SBLAS_ADVISE_SET_READ_MOSTLY: a = cudaMemAdviseSetReadMostly; break;
    case SBLAS_ADVISE_UNSET_READ_MOSTLY: a = cudaMemAdviseUnsetReadMostly; break;
    case SBLAS_ADVISE_SET_PREFERRED_LOCATION: a = cudaMemAdviseSetPreferredLocation; break;
    case SBLAS_ADVISE_UNSET_PREFERRED_LOCATION: a = cudaMemAdviseUnsetPreferredLocation; break;
    case SBLAS_ADVISE_SET_ACCESSED_BY: a = cudaMemAdviseSetAccessedBy; break;
    case SBLAS_ADVISE_UNSET_ACCESSED_BY: a = cudaMemAdviseUnsetAccessedBy; break;
    default: set_error("mem_advise: unknown advice %d", advice); return 2;
  }
  const int dev = gpu < 0 ? cudaCpuDeviceId : sblas_device_of(gpu);
  SBLAS_CUDA(cudaMemAdvise(p, bytes, a, dev));
  return 0;
}
int sblas_mem_prefetch(void* p, size_t bytes, int gpu) {
  if (!p || bytes == 0) return 0;
  const int dev = gpu < 0 ? cudaCpuDeviceId : sblas_device_of(gpu);
  if (gpu >= 0) {
    SBLAS_ON_GPU(gpu);
    SBLAS_CUDA(cudaMemPrefetchAsync(p, bytes, dev, nullptr));
  } else {
    SBLAS_CUDA(cudaMemPrefetchAsync(p, bytes, dev, nullptr));
  }
  return 0;
}

}  // extern "C"
