// convert.cu — SURVEY §8(f) f3: the format containers beside CSR (matrix.h:40-272, :457-506) on the
// device: CSR <-> CSC conversion and the COO (row, col) sort.  Not on the SpMV/SpMM hot path.
//
// The reference converts on the CPU with a histogram, a serial scan and a serial scatter in row
// order (matrix.h:40-64, :67-91); the result is the STABLE counting sort of the entries by
// column: inside a column the entries keep CSR order (rows ascending, duplicates in storage
// order).  The same result on the device, deterministically (an atomic scatter would not keep the
// order):
//   1. a stable LSD radix sort of (column, entry index) pairs — the one library piece, CUB's
//      DeviceRadixSort from the CUDA toolkit, over only the bits the column range needs;
//   2. colptr[c] = lower_bound(sorted columns, c), one thread per column (no atomics, no scan);
//   3. one gather pass: row of the entry by upper_bound in the row pointers, value by index.
// CSC -> CSR is the same routine with the roles of rows and columns exchanged.  The COO sort
// (CooSparseMatrix::sortByRow, matrix.h:221-237: qsort by (row, col)) is a stable radix sort of the
// 64-bit key row<<32|col, so equal (row, col) pairs keep file order, which qsort leaves open.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "sblas_b200.h"

namespace sblas {

__global__ void iota_kernel(int32_t* __restrict__ p, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = (int32_t)i;
}

// ptr[c] = number of sorted keys < c, c in [0, n]
__global__ void ptr_from_sorted_kernel(const int32_t* __restrict__ keys, int64_t nnz, int32_t n,
                                       int32_t* __restrict__ ptr) {
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c <= n; c += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = nnz;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < (int32_t)c) lo = mid + 1; else hi = mid;
    }
    ptr[c] = (int32_t)lo;
  }
}

// out_idx[i] = major index (row of CSR) of entry perm[i]; out_val[i] = val[perm[i]]
template <typename T>
__global__ void gather_major_kernel(const int32_t* __restrict__ perm, int64_t nnz, const int32_t* __restrict__ majptr,
                                    int32_t m, const T* __restrict__ val, int32_t* __restrict__ out_idx,
                                    T* __restrict__ out_val) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t j = perm[i];
    int32_t lo = 0, hi = m + 1;  // upper_bound(majptr[0..m], j) - 1: the non-empty row holding entry j
    while (lo < hi) {
      const int32_t mid = (lo + hi) >> 1;
      if (majptr[mid] <= j) lo = mid + 1; else hi = mid;
    }
    out_idx[i] = lo - 1;
    out_val[i] = val[j];
  }
}

__global__ void coo_keys_kernel(const int32_t* __restrict__ r, const int32_t* __restrict__ c, int64_t nnz,
                                uint64_t* __restrict__ keys, int32_t* __restrict__ idx) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    keys[i] = ((uint64_t)(uint32_t)r[i] << 32) | (uint32_t)c[i];
    idx[i] = (int32_t)i;
  }
}
template <typename T>
__global__ void coo_gather_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ perm, int64_t nnz,
                                  const T* __restrict__ val, int32_t* __restrict__ r, int32_t* __restrict__ c,
                                  T* __restrict__ out_val) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    r[i] = (int32_t)(keys[i] >> 32);
    c[i] = (int32_t)(keys[i] & 0xffffffffu);
    out_val[i] = val[perm[i]];
  }
}

static int bits_for(int64_t n) {  // bits needed for keys in [0, n)
  int b = 1;
  while (b < 63 && ((int64_t)1 << b) < n) ++b;
  return b;
}
static int grid_for(int64_t n) { return (int)std::min<int64_t>(std::max<int64_t>(ceil_div(n, 256), 1), 148 * 16); }

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) {
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
    if (e != cudaSuccess) { set_error("convert: cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); p = nullptr; return 1; }
    return 0;
  }
};

// minor-sorted copy of a compressed matrix: (majptr[m+1], minidx[nnz], val) -> (outptr[n+1], outidx, outval)
template <typename T>
static int transpose_compressed(cudaStream_t st, int32_t m, int32_t n, int64_t nnz, const int32_t* majptr,
                                const int32_t* minidx, const T* val, int32_t* outptr, int32_t* outidx, T* outval) {
  SBLAS_CHECK(m >= 0 && n >= 0 && nnz >= 0 && nnz < (1ll << 31), "convert: bad sizes");
  SBLAS_CHECK(outptr && (nnz == 0 || (majptr && minidx && val && outidx && outval)), "convert: null pointer");
  if (nnz == 0) {
    SBLAS_CUDA(cudaMemsetAsync(outptr, 0, sizeof(int32_t) * ((size_t)n + 1), st));
    return 0;
  }
  DevBuf keys, idx_in, idx_out, tmp;
  if (keys.alloc(sizeof(int32_t) * (size_t)nnz) || idx_in.alloc(sizeof(int32_t) * (size_t)nnz) ||
      idx_out.alloc(sizeof(int32_t) * (size_t)nnz))
    return 1;
  iota_kernel<<<grid_for(nnz), 256, 0, st>>>((int32_t*)idx_in.p, nnz);
  SBLAS_LAUNCH_CHECK();
  size_t tb = 0;
  const int eb = bits_for(std::max<int64_t>(n, 2));
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, minidx, (int32_t*)keys.p, (const int32_t*)idx_in.p,
                                             (int32_t*)idx_out.p, (int)nnz, 0, eb, st));
  if (tmp.alloc(tb)) return 1;
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tb, minidx, (int32_t*)keys.p, (const int32_t*)idx_in.p,
                                             (int32_t*)idx_out.p, (int)nnz, 0, eb, st));
  ptr_from_sorted_kernel<<<grid_for((int64_t)n + 1), 256, 0, st>>>((const int32_t*)keys.p, nnz, n, outptr);
  SBLAS_LAUNCH_CHECK();
  gather_major_kernel<T><<<grid_for(nnz), 256, 0, st>>>((const int32_t*)idx_out.p, nnz, majptr, m, val, outidx, outval);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaStreamSynchronize(st));  // the temporaries are released on return
  return 0;
}

template <typename T>
static int coo_sort(cudaStream_t st, int64_t nnz, int32_t height, int32_t* r, int32_t* c, T* val) {
  SBLAS_CHECK(nnz >= 0 && nnz < (1ll << 31) && height >= 0, "coo_sort: bad sizes");
  if (nnz == 0) return 0;
  SBLAS_CHECK(r && c && val, "coo_sort: null pointer");
  DevBuf k_in, k_out, i_in, i_out, v_in, tmp;
  if (k_in.alloc(8 * (size_t)nnz) || k_out.alloc(8 * (size_t)nnz) || i_in.alloc(4 * (size_t)nnz) ||
      i_out.alloc(4 * (size_t)nnz) || v_in.alloc(sizeof(T) * (size_t)nnz))
    return 1;
  coo_keys_kernel<<<grid_for(nnz), 256, 0, st>>>(r, c, nnz, (uint64_t*)k_in.p, (int32_t*)i_in.p);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaMemcpyAsync(v_in.p, val, sizeof(T) * (size_t)nnz, cudaMemcpyDeviceToDevice, st));
  size_t tb = 0;
  const int eb = 32 + bits_for(std::max<int64_t>(height, 2));
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, (const uint64_t*)k_in.p, (uint64_t*)k_out.p,
                                             (const int32_t*)i_in.p, (int32_t*)i_out.p, (int)nnz, 0, eb, st));
  if (tmp.alloc(tb)) return 1;
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tb, (const uint64_t*)k_in.p, (uint64_t*)k_out.p,
                                             (const int32_t*)i_in.p, (int32_t*)i_out.p, (int)nnz, 0, eb, st));
  coo_gather_kernel<T><<<grid_for(nnz), 256, 0, st>>>((const uint64_t*)k_out.p, (const int32_t*)i_out.p, nnz,
                                                      (const T*)v_in.p, r, c, val);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace sblas

using namespace sblas;

extern "C" {
int sblas_csr2csc_f64(void* stream, int32_t m, int32_t n, int64_t nnz, const int32_t* d_rowptr, const int32_t* d_colidx,
                      const double* d_val, int32_t* d_cscColPtr, int32_t* d_cscRowIdx, double* d_cscVal) {
  return transpose_compressed<double>((cudaStream_t)stream, m, n, nnz, d_rowptr, d_colidx, d_val, d_cscColPtr, d_cscRowIdx, d_cscVal);
}
int sblas_csr2csc_f32(void* stream, int32_t m, int32_t n, int64_t nnz, const int32_t* d_rowptr, const int32_t* d_colidx,
                      const float* d_val, int32_t* d_cscColPtr, int32_t* d_cscRowIdx, float* d_cscVal) {
  return transpose_compressed<float>((cudaStream_t)stream, m, n, nnz, d_rowptr, d_colidx, d_val, d_cscColPtr, d_cscRowIdx, d_cscVal);
}
int sblas_csc2csr_f64(void* stream, int32_t n, int32_t m, int64_t nnz, const int32_t* d_colptr, const int32_t* d_rowidx,
                      const double* d_val, int32_t* d_csrRowPtr, int32_t* d_csrColIdx, double* d_csrVal) {
  return transpose_compressed<double>((cudaStream_t)stream, n, m, nnz, d_colptr, d_rowidx, d_val, d_csrRowPtr, d_csrColIdx, d_csrVal);
}
int sblas_csc2csr_f32(void* stream, int32_t n, int32_t m, int64_t nnz, const int32_t* d_colptr, const int32_t* d_rowidx,
                      const float* d_val, int32_t* d_csrRowPtr, int32_t* d_csrColIdx, float* d_csrVal) {
  return transpose_compressed<float>((cudaStream_t)stream, n, m, nnz, d_colptr, d_rowidx, d_val, d_csrRowPtr, d_csrColIdx, d_csrVal);
}
int sblas_coo_sort_f64(void* stream, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col, double* d_val) {
  return coo_sort<double>((cudaStream_t)stream, nnz, height, d_row, d_col, d_val);
}
int sblas_coo_sort_f32(void* stream, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col, float* d_val) {
  return coo_sort<float>((cudaStream_t)stream, nnz, height, d_row, d_col, d_val);
}
}  // extern "C"
