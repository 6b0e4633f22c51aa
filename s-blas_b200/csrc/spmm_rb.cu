// spmm_rb.cu — row-block (R x 1 BCSR) device mirror of A and the SpMM kernel that runs on it.
//
// Why: with the CSR kernel of spmm.cu every nnz fetches its k-wide B row through L1 once, so
// structured inputs (banded / stencil / FEM: adjacent rows share most columns) are bound by the
// L1 data path (128 B/clk/SM: 7.6 ms for BASELINE config 5) long before HBM (2.2 ms).  The only
// way under that bound is register reuse: ONE load of a B row slice feeding the accumulators of
// the R adjacent A rows that have a nonzero in that column.  That is a sparse format question,
// and "matrix.h (device CSR layout and padding)" is one of the subsystems the north star lets
// change: CsrSparseMatrix::sync2gpu builds, next to the CSR arrays the reference's API exposes,
// a mirror in which R consecutive rows form a block row and every distinct column of a block
// row is one block = { col, presence mask, R values }.  The mirror is only kept when it pays
// (fill = nnz / (blocks * R) >= 1/2); uniform-random inputs stay on the CSR kernel.
//
// Arithmetic: row r of a block row accumulates ITS nonzeros in CSR storage order (the R-way
// merge consumes every row front to back), with separately rounded multiply and add unless
// SBLAS_FLAG_FMA is set — so results stay bit-identical to sblas_spmm_csr_cpu (spmm.h:56-68),
// absent entries are skipped through the mask (no 0 * Inf).
#include <cub/cub.cuh>

#include "plan.cuh"

namespace sblas {

// ---- build: R-way merge of the rows of every block row -----------------------------------------
template <int R, bool FILL, typename T>
__global__ void __launch_bounds__(128)
rb_merge_kernel(int32_t rows, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                const T* __restrict__ val, int32_t* __restrict__ counts,
                const int32_t* __restrict__ blk_ptr, int2* __restrict__ colmask, T* __restrict__ bval) {
  const int64_t br = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nbr = ((int64_t)rows + R - 1) / R;
  if (br >= nbr) return;
  int32_t p[R], e[R], head[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t row = br * R + r;
    p[r] = row < rows ? rowptr[row] : 0;
    e[r] = row < rows ? rowptr[row + 1] : 0;
    head[r] = p[r] < e[r] ? colidx[p[r]] : INT32_MAX;
  }
  int32_t n = 0;
  int64_t out = FILL ? blk_ptr[br] : 0;
  while (true) {
    int32_t c = head[0];
#pragma unroll
    for (int r = 1; r < R; ++r) c = min(c, head[r]);
    if (c == INT32_MAX) break;
    uint32_t mask = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (head[r] == c) {
        mask |= 1u << r;
        if (FILL) bval[(out + n) * R + r] = val[p[r]];
        ++p[r];
        head[r] = p[r] < e[r] ? colidx[p[r]] : INT32_MAX;
      } else if (FILL) {
        bval[(out + n) * R + r] = T(0);
      }
    }
    if (FILL) colmask[out + n] = make_int2(c, (int)mask);
    ++n;
  }
  if (!FILL) counts[br] = n;
}

// most blocks in W consecutive block rows (staging capacity of the pipelined kernel)
__global__ void rb_tile_max_kernel(const int32_t* __restrict__ blk_ptr, int64_t n_blockrows, int W, int32_t* out) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t n_tiles = (n_blockrows + W - 1) / W;
  int32_t v = 0;
  if (t < n_tiles) v = blk_ptr[min((t + 1) * W, n_blockrows)] - blk_ptr[t * W];
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) atomicMax(out, v);
}

// distinct columns of every tile of W = 8 block rows: 8-way merge of the block rows' (ascending)
// column lists.  Pass 1 (FILL = false) counts and checks that every block row IS ascending
// (*unsorted is raised otherwise); pass 2 writes the list and tags every block with its position.
template <bool FILL>
__global__ void __launch_bounds__(128)
rb_tile_cols_kernel(int64_t n_blockrows, const int32_t* __restrict__ blk_ptr, int2* __restrict__ colmask,
                    int32_t* __restrict__ counts, const int32_t* __restrict__ colptr,
                    int32_t* __restrict__ tile_cols, int32_t* __restrict__ unsorted) {
  constexpr int W = 8;
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t n_tiles = (n_blockrows + W - 1) / W;
  if (t >= n_tiles) return;
  int32_t p[W], e[W], head[W];
#pragma unroll
  for (int j = 0; j < W; ++j) {
    const int64_t br = t * W + j;
    p[j] = br < n_blockrows ? blk_ptr[br] : 0;
    e[j] = br < n_blockrows ? blk_ptr[br + 1] : 0;
    head[j] = p[j] < e[j] ? colmask[p[j]].x : INT32_MAX;
  }
  int32_t n = 0;
  bool bad = false;
  const int64_t out = FILL ? colptr[t] : 0;
  while (true) {
    int32_t c = head[0];
#pragma unroll
    for (int j = 1; j < W; ++j) c = min(c, head[j]);
    if (c == INT32_MAX) break;
#pragma unroll
    for (int j = 0; j < W; ++j) {
      if (head[j] == c) {
        if (FILL) colmask[p[j]].y = (colmask[p[j]].y & kRbMaskBits) | (n << 9);
        ++p[j];
        const int32_t nh = p[j] < e[j] ? colmask[p[j]].x : INT32_MAX;
        bad = bad || nh <= c;
        head[j] = nh;
      }
    }
    if (FILL) tile_cols[out + n] = c;
    ++n;
    if (bad) break;
  }
  if (!FILL) {
    counts[t] = n;
    if (bad) atomicExch(unsorted, 1);
  }
}

// builds rb.d_tile_colptr / d_tile_cols / max_tile_cols (left empty when some row is not ascending)
static int build_tile_cols(cudaStream_t st, RowBlocks& rb) {
  const int64_t n_tiles = ceil_div(rb.n_blockrows, 8);
  int32_t* counts = nullptr;
  SBLAS_CUDA(cudaMalloc(&counts, sizeof(int32_t) * (size_t)(n_tiles + 2)));
  SBLAS_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * (size_t)(n_tiles + 2), st));
  int32_t* flag = counts + n_tiles + 1;
  const int grid = (int)ceil_div(n_tiles, 128);
  rb_tile_cols_kernel<false><<<grid, 128, 0, st>>>(rb.n_blockrows, rb.d_ptr, (int2*)rb.d_colmask, counts, nullptr,
                                                   nullptr, flag);
  SBLAS_LAUNCH_CHECK();
  int32_t unsorted = 0;
  SBLAS_CUDA(cudaMemcpyAsync(&unsorted, flag, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  if (unsorted) {
    SBLAS_CUDA(cudaFree(counts));
    return 0;
  }
  SBLAS_CUDA(cudaMalloc(&rb.d_tile_colptr, sizeof(int32_t) * (size_t)(n_tiles + 1)));
  void* tmp = nullptr;
  size_t tb = 0;
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, counts, rb.d_tile_colptr, (int)(n_tiles + 1), st));
  SBLAS_CUDA(cudaMalloc(&tmp, tb));
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, counts, rb.d_tile_colptr, (int)(n_tiles + 1), st));
  void* mx = nullptr;
  size_t mb = 0;
  int32_t* d_max = nullptr;
  SBLAS_CUDA(cudaMalloc(&d_max, sizeof(int32_t)));
  SBLAS_CUDA(cub::DeviceReduce::Max(nullptr, mb, counts, d_max, (int)n_tiles, st));
  SBLAS_CUDA(cudaMalloc(&mx, mb));
  SBLAS_CUDA(cub::DeviceReduce::Max(mx, mb, counts, d_max, (int)n_tiles, st));
  int32_t total = 0, max_cols = 0;
  SBLAS_CUDA(cudaMemcpyAsync(&total, rb.d_tile_colptr + n_tiles, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaMemcpyAsync(&max_cols, d_max, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  SBLAS_CUDA(cudaFree(tmp));
  SBLAS_CUDA(cudaFree(mx));
  SBLAS_CUDA(cudaFree(d_max));
  SBLAS_CUDA(cudaMalloc(&rb.d_tile_cols, sizeof(int32_t) * (size_t)std::max(total, 1)));
  rb_tile_cols_kernel<true><<<grid, 128, 0, st>>>(rb.n_blockrows, rb.d_ptr, (int2*)rb.d_colmask, nullptr,
                                                  rb.d_tile_colptr, rb.d_tile_cols, nullptr);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaStreamSynchronize(st));
  SBLAS_CUDA(cudaFree(counts));
  rb.max_tile_cols = max_cols;
  return 0;
}

template <typename T, int R>
static int build_rowblocks(cudaStream_t st, sblasPlan_st* plan, const int32_t* ci, const T* val, bool force,
                           double min_fill) {
  const int64_t nbr = ((int64_t)plan->rows + R - 1) / R;
  RowBlocks& rb = plan->rb;
  int32_t* counts = nullptr;
  SBLAS_CUDA(cudaMalloc(&counts, sizeof(int32_t) * (size_t)(nbr + 4)));  // >= 3 slots for the tile statistics
  SBLAS_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * (size_t)(nbr + 1), st));
  const int grid = (int)ceil_div(nbr, 128);
  rb_merge_kernel<R, false, T><<<grid, 128, 0, st>>>(plan->rows, plan->d_rowptr, ci, val, counts, nullptr, nullptr, nullptr);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaMalloc(&rb.d_ptr, sizeof(int32_t) * (size_t)(nbr + 1)));
  void* tmp = nullptr;
  size_t tb = 0;
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, counts, rb.d_ptr, (int)(nbr + 1), st));
  SBLAS_CUDA(cudaMalloc(&tmp, tb));
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, counts, rb.d_ptr, (int)(nbr + 1), st));
  int32_t total = 0;
  SBLAS_CUDA(cudaMemcpyAsync(&total, rb.d_ptr + nbr, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  SBLAS_CUDA(cudaFree(tmp));
  rb.R = R;
  rb.n_blockrows = nbr;
  rb.n_blocks = total;
  rb.is_f64 = sizeof(T) == 8;
  rb.fill = total > 0 ? (double)plan->nnz / ((double)total * R) : 0.0;
  if (rb.fill < min_fill && !force) {  // does not pay: keep only the statistic
    SBLAS_CUDA(cudaFree(counts));
    SBLAS_CUDA(cudaFree(rb.d_ptr));
    rb.d_ptr = nullptr;
    rb.n_blocks = 0;
    return 0;
  }
  SBLAS_CUDA(cudaMalloc(&rb.d_colmask, sizeof(int2) * (size_t)std::max(total, 1)));
  SBLAS_CUDA(cudaMalloc(&rb.d_val, sizeof(T) * (size_t)std::max(total, 1) * R));
  rb_merge_kernel<R, true, T><<<grid, 128, 0, st>>>(plan->rows, plan->d_rowptr, ci, val, counts, rb.d_ptr,
                                                    (int2*)rb.d_colmask, (T*)rb.d_val);
  SBLAS_LAUNCH_CHECK();
  // counts[0..2] are free again: reuse them for the tile statistics (W = 8, 4, 2 block rows)
  SBLAS_CUDA(cudaMemsetAsync(counts, 0, 3 * sizeof(int32_t), st));
  int32_t tile_max[3] = {0, 0, 0};
  for (int j = 0; j < 3; ++j) {
    const int W = 8 >> j;
    const int64_t n_tiles = ceil_div(nbr, W);
    rb_tile_max_kernel<<<(int)ceil_div(n_tiles, 256), 256, 0, st>>>(rb.d_ptr, nbr, W, counts + j);
    SBLAS_LAUNCH_CHECK();
  }
  SBLAS_CUDA(cudaMemcpyAsync(tile_max, counts, 3 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  rb.max_tile_blocks = tile_max[0];
  rb.max_tile_blocks4 = tile_max[1];
  rb.max_tile_blocks2 = tile_max[2];
  SBLAS_CUDA(cudaFree(counts));
  return build_tile_cols(st, rb);
}

// R = 0 (what sync2gpu asks for): R = 4, kept when its fill reaches kRowBlockMinFill.  R = 4 / 8:
// that height or nothing (R = 8 needs fill >= kRowBlockMinFill8; measured no faster than R = 4
// on the banded inputs — 9.9 against 8.8 ms on b32m, 8.0 against 7.3 ms on config 5 — so it is
// never chosen automatically).  R < 0: |R|, kept whatever the fill (tests).
template <typename T>
static int build_impl(cudaStream_t st, sblasPlan_st* plan, const int32_t* ci, const T* val, int32_t R) {
  SBLAS_CHECK(plan && (plan->nnz == 0 || (ci && val)), "plan_build_rowblocks: null argument");
  rowblocks_free(plan);
  if (plan->rows == 0 || plan->nnz == 0) return 0;
  const bool force = R < 0;
  if (R < 0) R = -R;
  SBLAS_CHECK(R == 0 || R == 4 || R == 8, "plan_build_rowblocks: R must be 0 (auto), 4 or 8");
  if (R == 8) return build_rowblocks<T, 8>(st, plan, ci, val, force, kRowBlockMinFill8);
  if (R == 4) return build_rowblocks<T, 4>(st, plan, ci, val, force, kRowBlockMinFill);
  return build_rowblocks<T, 4>(st, plan, ci, val, false, kRowBlockMinFill);
}

void rowblocks_free(sblasPlan_st* plan) {
  RowBlocks& rb = plan->rb;
  if (rb.d_ptr) cudaFree(rb.d_ptr);
  if (rb.d_colmask) cudaFree(rb.d_colmask);
  if (rb.d_val) cudaFree(rb.d_val);
  if (rb.d_tile_colptr) cudaFree(rb.d_tile_colptr);
  if (rb.d_tile_cols) cudaFree(rb.d_tile_cols);
  rb = RowBlocks();
}

}  // namespace sblas

using namespace sblas;
extern "C" {
int sblas_plan_build_rowblocks_f64(void* stream, sblasPlan_t plan, const int32_t* d_colidx, const double* d_val, int32_t R) {
  return build_impl<double>((cudaStream_t)stream, plan, d_colidx, d_val, R);
}
int sblas_plan_build_rowblocks_f32(void* stream, sblasPlan_t plan, const int32_t* d_colidx, const float* d_val, int32_t R) {
  return build_impl<float>((cudaStream_t)stream, plan, d_colidx, d_val, R);
}
int sblas_plan_rowblock_stats(sblasPlan_t plan, int32_t* R, int64_t* n_blocks, double* fill) {
  SBLAS_CHECK(plan, "plan_rowblock_stats: null plan");
  if (R) *R = plan->rb.R;
  if (n_blocks) *n_blocks = plan->rb.n_blocks;
  if (fill) *fill = plan->rb.fill;
  return 0;
}
}
