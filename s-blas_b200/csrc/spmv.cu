// spmv.cu — CSR SpMV kernels for sm_100a:  y <- beta*y + alpha*A*x.
//
// Replaces cusparseSpMV (spmv.h:96-106) and the separate axpby pass (spmv.h:138,
// kernel.h:28-37) of the reference; the CPU statement of the same arithmetic is
// sblas_spmv_csr_cpu (spmv.h:16-32).
//
// Variants (sblasOpts.variant; AUTO picks from the plan's row statistics and mirrors).  Since round 2 the first
// choice is SELL / COLPANEL (spmv_sell_kernel below + colpanel.cu): a lane per row over the sliced-ELL mirror that
// sync2gpu builds — one panel for regular inputs, column panels with L2-resident x slices for gather-heavy ones, the
// virtual-row form for skewed ones.  Without a mirror:
//
//  TILE   nnz-balanced, row-aligned merge-path tiles (plan.cuh).  A CTA streams its tile's
//         col_idx / val with 128-bit L1-bypassing loads (fully coalesced whatever the row
//         lengths), gathers x through L1, and parks the products in shared memory; then
//         rows are reduced from shared memory by 1 lane (short rows; sequential order,
//         hence bit-identical to the reference CPU loop) or by a 2..32-lane group with a
//         warp-shuffle reduction.  The beta/alpha epilogue and the stores to every GPU's
//         replica of y are fused.   HBM-bound; algorithmic bytes per nnz = sizeof(T)+4,
//         per row = 4 + 2*sizeof(T)  (+ x once).
//  VECTOR classic sub-warp-per-row kernel straight from global memory; needs no plan.
//  SPLIT  merge-path tiles in nnz space that cut long rows; partial sums of cut rows go
//         to a carry array and the last CTA to finish folds them in (skewed inputs: R-MAT).
#include <mutex>

#include "plan.cuh"

namespace sblas {

constexpr int kSpmvBlock = 256;
constexpr int kSpmvDefaultTile = 3072;      // TILE
constexpr int kSpmvSplitDefaultTile = 1536; // SPLIT: config 4 0.66 ms (1536-2048) against 1.5 ms at 3072 and up
constexpr int kSpmvPipeDefaultTile = 1536;  // PIPE: small tiles, 2 stages -> most resident CTAs (sweeps, profiles/)
constexpr int kSpmvMaxSmem = 96 * 1024;
constexpr int kSpmvPipeMaxRow = 4096;       // longest row PIPE stages whole (2 stages <= ~200 KB)

template <typename T>
struct RowSink {
  T* y;
  OutSpec<T> out;
  T* carry;       // head partial sum at [0], tail partial sum at [4] (= kpad of k = 1, sblas_b200.h)
  int32_t rows;   // local row count
  int32_t flags;  // HEAD_PARTIAL | TAIL_PARTIAL
  T alpha, beta;
  // column-panelled SpMV (colpanel.cu): the passes over the panels of A carry the RAW running row sums in
  // `part`.  phase 0 = plain call; 1 = first panel (part = sum); 2 = middle panel (part = part (+) sum, the
  // accumulation CONTINUES from the carried value); 3 = last panel (y = beta*y + alpha*(part (+) sum)).
  // stage_src = what the PIPE kernel stages per row beside the tile: y (phase 0), nothing (1), part (2, 3).
  T* part;
  const T* stage_src;
  int32_t phase;

  __device__ __forceinline__ void put(int32_t r, T sum) const {
    if ((flags & SBLAS_FLAG_HEAD_PARTIAL) && r == 0) { carry[0] = sum; return; }
    if ((flags & SBLAS_FLAG_TAIL_PARTIAL) && r == rows - 1) { carry[4] = sum; return; }
    T res = axpby_rn(alpha, sum, beta, y[r]);
    y[r] = res;
    for (int d = 0; d < out.n; ++d) out.ptr[d][r] = res;
  }
  // same, with the row's staged value (old y[r]; in phases 2 / 3 the carried sum, which `sum` already continues)
  __device__ __forceinline__ void put_y(int32_t r, T sum, T staged) const {
    if (phase == 1 || phase == 2) { part[r] = sum; return; }
    if ((flags & SBLAS_FLAG_HEAD_PARTIAL) && r == 0) { carry[0] = sum; return; }
    if ((flags & SBLAS_FLAG_TAIL_PARTIAL) && r == rows - 1) { carry[4] = sum; return; }
    T res = axpby_rn(alpha, sum, beta, phase == 3 ? y[r] : staged);
    y[r] = res;
    for (int d = 0; d < out.n; ++d) out.ptr[d][r] = res;
  }
};

// ---- 4-element helpers -------------------------------------------------------------------
__device__ __forceinline__ void load4_stream(const double* val, int g, double (&v)[4]) {
  double2 a = ldg_stream_d2(reinterpret_cast<const double2*>(val) + 2 * (size_t)g);
  double2 b = ldg_stream_d2(reinterpret_cast<const double2*>(val) + 2 * (size_t)g + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void load4_stream(const float* val, int g, float (&v)[4]) {
  float4 a = ldg_stream_f4(reinterpret_cast<const float4*>(val) + (size_t)g);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void store4_smem(double* s, const double (&p)[4]) {
  reinterpret_cast<double2*>(s)[0] = make_double2(p[0], p[1]);
  reinterpret_cast<double2*>(s)[1] = make_double2(p[2], p[3]);
}
__device__ __forceinline__ void store4_smem(float* s, const float (&p)[4]) {
  reinterpret_cast<float4*>(s)[0] = make_float4(p[0], p[1], p[2], p[3]);
}

// Reduce the rows of a tile from shared-memory products with G lanes per row.
template <typename T, int G, int BLOCK, typename Sink>
__device__ __forceinline__ void reduce_rows(const T* s_prod, const int32_t* s_rp, int32_t nrows,
                                            int32_t n0a, const Sink& sink) {
  const int tid = threadIdx.x;
  if (G == 1) {
    for (int i = tid; i < nrows; i += BLOCK) {
      const int s = s_rp[i] - n0a, e = s_rp[i + 1] - n0a;
      T sum = 0;
      for (int j = s; j < e; ++j) sum = add_rn(sum, s_prod[j]);  // order of spmv.h:24-30
      sink.put(i, sum);
    }
  } else {
    const int gl = tid & (G - 1), gi = tid / G;
    for (int i0 = 0; i0 < nrows; i0 += BLOCK / G) {  // block-uniform trip count
      const int i = i0 + gi;
      const bool valid = i < nrows;
      const int s = valid ? s_rp[i] - n0a : 0, e = valid ? s_rp[i + 1] - n0a : 0;
      T sum = 0;
      for (int j = s + gl; j < e; j += G) sum += s_prod[j];
#pragma unroll
      for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
      if (valid && gl == 0) sink.put(i, sum);
    }
  }
}

template <typename T, int BLOCK, typename Sink>
__device__ __forceinline__ void reduce_rows_dispatch(int lanes, const T* s_prod,
                                                     const int32_t* s_rp, int32_t nrows,
                                                     int32_t n0a, const Sink& sink) {
  switch (lanes) {
    case 1: reduce_rows<T, 1, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
    case 2: reduce_rows<T, 2, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
    case 4: reduce_rows<T, 4, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
    case 8: reduce_rows<T, 8, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
    case 16: reduce_rows<T, 16, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
    default: reduce_rows<T, 32, BLOCK>(s_prod, s_rp, nrows, n0a, sink); break;
  }
}

// Same, for SPLIT tiles: row pointers are clamped to the tile's nnz window [lo, hi).
template <typename T, int G, int BLOCK, typename Sink>
__device__ __forceinline__ void reduce_rows_clamped(const T* s_prod, const int32_t* s_rp, int32_t nrows,
                                                    int32_t n0a, int32_t lo, int32_t hi, const Sink& sink) {
  const int tid = threadIdx.x;
  const int gl = tid & (G - 1), gi = tid / G;
  for (int i0 = 0; i0 < nrows; i0 += BLOCK / G) {  // block-uniform trip count
    const int i = i0 + gi;
    const bool valid = i < nrows;
    const int s = valid ? min(max(s_rp[i], lo), hi) - n0a : 0;
    const int e = valid ? min(max(s_rp[i + 1], lo), hi) - n0a : 0;
    T sum = 0;
    if (G == 1) {
      for (int j = s; j < e; ++j) sum = add_rn(sum, s_prod[j]);
    } else {
      for (int j = s + gl; j < e; j += G) sum += s_prod[j];
#pragma unroll
      for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    }
    if (valid && gl == 0) sink.put(i, sum);
  }
}
template <typename T, int BLOCK, typename Sink>
__device__ __forceinline__ void reduce_rows_clamped_dispatch(int lanes, const T* s_prod, const int32_t* s_rp,
                                                             int32_t nrows, int32_t n0a, int32_t lo, int32_t hi,
                                                             const Sink& sink) {
  switch (lanes) {
    case 1: reduce_rows_clamped<T, 1, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
    case 2: reduce_rows_clamped<T, 2, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
    case 4: reduce_rows_clamped<T, 4, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
    case 8: reduce_rows_clamped<T, 8, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
    case 16: reduce_rows_clamped<T, 16, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
    default: reduce_rows_clamped<T, 32, BLOCK>(s_prod, s_rp, nrows, n0a, lo, hi, sink); break;
  }
}

__device__ __forceinline__ int lanes_for_avg(int avg) {
  return avg <= 6 ? 1 : avg <= 12 ? 2 : avg <= 24 ? 4 : avg <= 48 ? 8 : avg <= 96 ? 16 : 32;
}

template <typename T>
struct TileSink {
  const RowSink<T>& sink;
  int32_t r0;
  __device__ __forceinline__ void put(int i, T sum) const { sink.put(r0 + i, sum); }
};

// Phase 1 shared by TILE and SPLIT: products of nnz [n0, n1) -> s_prod[j - (n0 & ~3)].
// NOL1: gather x past L1 (ld.global.nc.L1::no_allocate) — an in-flight gather that allocates holds an L1
// line, and with most of the 256 KB carved out as shared memory the lines cap the gathers in flight.
template <typename T, int BLOCK, bool VEC, bool NOL1 = false>
__device__ __forceinline__ void stage_products(const int32_t* __restrict__ colidx,
                                               const T* __restrict__ val,
                                               const T* __restrict__ x, int32_t n0, int32_t n1,
                                               T* s_prod) {
  auto gx = [&](int32_t c) -> T { return NOL1 ? ldg_stream(x + c) : __ldg(x + c); };
  const int tid = threadIdx.x;
  const int32_t n0a = n0 & ~3;
  if (VEC) {
    const int g0 = n0a >> 2, g1 = (int)(((int64_t)n1 + 3) >> 2);
#pragma unroll 2
    for (int g = g0 + tid; g < g1; g += BLOCK) {
      const int base = g << 2;
      T p[4];
      if (base >= n0 && base + 4 <= n1) {
        const int4 c = ldg_stream_i4(reinterpret_cast<const int4*>(colidx) + g);
        T v[4];
        load4_stream(val, g, v);
        p[0] = mul_rn(v[0], gx(c.x));
        p[1] = mul_rn(v[1], gx(c.y));
        p[2] = mul_rn(v[2], gx(c.z));
        p[3] = mul_rn(v[3], gx(c.w));
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = base + e;
          p[e] = (j >= n0 && j < n1) ? mul_rn(ldg_stream(val + j), gx(ldg_stream(colidx + j)))
                                     : T(0);
        }
      }
      store4_smem(s_prod + (base - n0a), p);
    }
  } else {
#pragma unroll 4
    for (int j = n0 + tid; j < n1; j += BLOCK)
      s_prod[j - n0a] = mul_rn(ldg_stream(val + j), gx(ldg_stream(colidx + j)));
  }
}

template <typename T, int BLOCK, bool VEC>
__global__ void __launch_bounds__(BLOCK)
spmv_tile_kernel(const int32_t* __restrict__ tile_row, const int32_t* __restrict__ rowptr,
                 const int32_t* __restrict__ colidx, const T* __restrict__ val,
                 const T* __restrict__ x, RowSink<T> sink, int32_t cap_nnz, int32_t force_lanes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* s_prod = reinterpret_cast<T*>(smem_raw);
  int32_t* s_rp = reinterpret_cast<int32_t*>(s_prod + cap_nnz);
  const int tid = threadIdx.x;
  const int32_t r0 = tile_row[blockIdx.x], r1 = tile_row[blockIdx.x + 1];
  const int32_t nrows = r1 - r0;
  if (nrows <= 0) return;
  for (int i = tid; i <= nrows; i += BLOCK) s_rp[i] = rowptr[r0 + i];
  __syncthreads();
  const int32_t n0 = s_rp[0], n1 = s_rp[nrows];
  const int32_t n0a = n0 & ~3;  // products are parked at s_prod[j - n0a]: 16-byte aligned groups

  // ---- phase 1: stream A, gather x, park products
  stage_products<T, BLOCK, VEC>(colidx, val, x, n0, n1, s_prod);
  __syncthreads();

  // ---- phase 2: rows from shared memory; lanes per row from the tile's mean row length
  int lanes = force_lanes;
  if (lanes == 0) {
    lanes = lanes_for_avg((n1 - n0) / nrows);
  }
  const TileSink<T> ts{sink, r0};
  reduce_rows_dispatch<T, BLOCK>(lanes, s_prod, s_rp, nrows, n0a, ts);
}

// ---- PIPE: persistent CTAs, NSTAGE-deep TMA-bulk staging of the tiles -------------------------
// The TILE kernel above pays three dependent DRAM round trips per CTA (row pointers -> A
// stream -> old y) with nothing in flight in between, so it sits near half of the HBM rate.
// Here a CTA walks the tiles b = blockIdx.x, blockIdx.x + gridDim.x, ... and keeps the next
// NSTAGE-1 tiles in flight while it reduces the current one.  Everything a tile needs is a
// contiguous range known from the plan's {tile_row, tile_nnz} tables:
//   col_idx[n0,n1), val[n0,n1)   one cp.async.bulk each for the 16-byte aligned middle
//                                 (issued by one thread), <= 3-element heads/tails by LDGSTS;
//   rowptr[r0..r1], y[r0,r1)     4/8-byte LDGSTS spread over the CTA;
// and ONE mbarrier phase per stage collects all of it (expect_tx for the bulk bytes,
// cp.async.mbarrier.arrive.noinc for the LDGSTS of every thread).  Products overwrite the
// staged values in place, rows are reduced from shared memory exactly as in TILE.
constexpr int kPipeHeaderBytes = 256;  // mbarriers + tile descriptors

template <typename T>
struct PipeStage {
  T* val;        // [cap_nnz]  staged values, then products (index j - (n0 & ~3))
  int32_t* col;  // [cap_nnz]
  int32_t* rp;   // [cap_rows + 1]
  T* y;          // [cap_rows]
};

template <typename T>
__host__ __device__ inline size_t pipe_stage_bytes(int32_t cap_nnz, int32_t cap_rows) {
  const size_t rp = ((size_t)(cap_rows + 1) * 4 + 15) & ~(size_t)15;
  const size_t y = ((size_t)cap_rows * sizeof(T) + 15) & ~(size_t)15;
  return (size_t)cap_nnz * (sizeof(T) + 4) + rp + y;  // cap_nnz is a multiple of 4
}

template <typename T>
__device__ __forceinline__ PipeStage<T> pipe_stage(unsigned char* base, int s, int32_t cap_nnz,
                                                   int32_t cap_rows) {
  unsigned char* p = base + (size_t)s * pipe_stage_bytes<T>(cap_nnz, cap_rows);
  PipeStage<T> st;
  st.val = reinterpret_cast<T*>(p);
  p += (size_t)cap_nnz * sizeof(T);
  st.col = reinterpret_cast<int32_t*>(p);
  p += (size_t)cap_nnz * 4;
  st.rp = reinterpret_cast<int32_t*>(p);
  p += ((size_t)(cap_rows + 1) * 4 + 15) & ~(size_t)15;
  st.y = reinterpret_cast<T*>(p);
  return st;
}

template <typename T>
struct PipeSink {
  const RowSink<T>& sink;
  const T* s_y;
  int32_t r0;
  __device__ __forceinline__ void put(int i, T sum) const { sink.put_y(r0 + i, sum, s_y[i]); }
  // value the row's accumulation starts from: the sum carried over from the previous column panels
  __device__ __forceinline__ T init(int i) const { return sink.phase >= 2 ? s_y[i] : T(0); }
};

// Rows of a staged tile: G lanes per row read (col, val) pairs from shared memory (consecutive
// lanes -> consecutive elements: conflict-free for G > 1; for G = 1 the stride is the row
// length), gather x with U independent loads in flight per lane, and accumulate in CSR order
// with separately rounded mul/add (G = 1: bit-identical to spmv.h:24-30), then shuffle-reduce.
template <typename T, int G, int U, int BLOCK, bool CLAMP, typename Sink>
__device__ __forceinline__ void consume_rows(const int32_t* __restrict__ s_col,
                                             const T* __restrict__ s_val,
                                             const int32_t* __restrict__ s_rp, int32_t nrows,
                                             int32_t n0a, int32_t lo, int32_t hi,
                                             const T* __restrict__ x, const Sink& sink) {
  const int gl = threadIdx.x & (G - 1), gi = threadIdx.x / G;
  for (int i0 = 0; i0 < nrows; i0 += BLOCK / G) {  // block-uniform trip count
    const int i = i0 + gi;
    const bool valid = i < nrows;
    int s = valid ? s_rp[i] : 0, e = valid ? s_rp[i + 1] : 0;
    if (CLAMP) {  // SPLIT tiles: rows are cut at the tile's nnz window [lo, hi)
      s = min(max(s, lo), hi);
      e = min(max(e, lo), hi);
    }
    s -= n0a;
    e -= n0a;
    T sum = (valid && gl == 0) ? sink.init(i) : T(0);
    for (int j0 = s + gl; j0 < e; j0 += G * U) {
      T vv[U], xx[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = j0 + u * G;
        if (j < e) {
          vv[u] = s_val[j];
          xx[u] = __ldg(x + s_col[j]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (j0 + u * G < e) sum = add_rn(sum, mul_rn(vv[u], xx[u]));
    }
    if (G > 1) {
#pragma unroll
      for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    }
    if (valid && gl == 0) sink.put(i, sum);
  }
}

// Short rows (the column-panel passes: ~8 nnz per row and panel), two phases on the staged tile.
// Why: with one lane per row the 8-element row stride makes the shared-memory reads of a warp 8-way (col_idx)
// and 16-way (8-byte values) bank conflicts; those wavefronts share the L1TEX data pipe with the x gathers
// (32 wavefronts per warp gather: every lane another line), 56 wavefronts per 32 nnz in all, and the pipe is
// what the pass is bound by.  Phase A is flat and balanced: thread t takes elements t, t + BLOCK, ... of the
// tile (conflict-free reads, UA independent gathers in flight per thread), the products overwrite the staged
// values.  Phase B: G consecutive lanes read the consecutive products of ONE row (conflict-free, shared-memory
// latency only) and shuffle-reduce; lane 0 adds the carried sum.  ~44 wavefronts per 32 nnz.
template <typename T, int G, int BLOCK, typename Sink>
__device__ __forceinline__ void consume_rows_flat(const int32_t* __restrict__ s_col, T* __restrict__ s_val,
                                                  const int32_t* __restrict__ s_rp, int32_t nrows,
                                                  int32_t n0, int32_t n1, const T* __restrict__ x,
                                                  const Sink& sink) {
  constexpr int UA = 6;
  const int tid = threadIdx.x;
  const int n0a = n0 & ~3, lo = n0 - n0a, hi = n1 - n0a;
  for (int j0 = lo + tid; j0 < hi; j0 += BLOCK * UA) {
    T vv[UA], xx[UA];
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int j = j0 + u * BLOCK;
      if (j < hi) {
        vv[u] = s_val[j];
        xx[u] = __ldg(x + s_col[j]);
      }
    }
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int j = j0 + u * BLOCK;
      if (j < hi) s_val[j] = mul_rn(vv[u], xx[u]);
    }
  }
  __syncthreads();
  constexpr int RPI = BLOCK / G;
  const int gl = tid & (G - 1), gi = tid / G;
  for (int i0 = 0; i0 < nrows; i0 += RPI) {  // block-uniform trip count
    const int i = i0 + gi;
    const bool valid = i < nrows;
    const int s = valid ? s_rp[i] - n0a : 0, e = valid ? s_rp[i + 1] - n0a : 0;
    T sum = 0;
    for (int j = s + gl; j < e; j += G) sum = add_rn(sum, s_val[j]);
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    if (valid && gl == 0) sink.put(i, add_rn(sink.init(i), sum));
  }
}

// A tile made of a few long row pieces (mean >= kPipeBlockModeAvg nnz): the whole CTA strides
// over one row at a time, warp-shuffle + shared-memory reduction, one barrier per row.
constexpr int kPipeBlockModeAvg = 128;
template <typename T, int BLOCK, typename Sink>
__device__ __forceinline__ void consume_rows_block(const int32_t* __restrict__ s_col,
                                                   const T* __restrict__ s_val,
                                                   const int32_t* __restrict__ s_rp, int32_t nrows,
                                                   int32_t n0a, int32_t lo, int32_t hi,
                                                   const T* __restrict__ x, const Sink& sink, T* red) {
  constexpr int NW = BLOCK / 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = 0; i < nrows; ++i) {
    const int s = min(max(s_rp[i], lo), hi) - n0a, e = min(max(s_rp[i + 1], lo), hi) - n0a;
    T sum = 0;
    for (int j0 = s + tid; j0 < e; j0 += BLOCK * 4) {
      T vv[4], xx[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + u * BLOCK;
        if (j < e) { vv[u] = s_val[j]; xx[u] = __ldg(x + s_col[j]); }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (j0 + u * BLOCK < e) sum = add_rn(sum, mul_rn(vv[u], xx[u]));
    }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    T* slot = red + (i & 1) * NW;  // two banks: one barrier per row is enough
    if (lane == 0) slot[warp] = sum;
    __syncthreads();
    if (tid == 0) {
      T tot = 0;
#pragma unroll
      for (int w = 0; w < NW; ++w) tot += slot[w];
      sink.put(i, tot);
    }
  }
}

// SPLIT flavour of the sink: the piece of a row cut by the end of the tile goes to carry_out, the
// last piece of a row begun in an earlier tile to head_val (folded by spmv_split_fixup_kernel)
template <typename T>
struct PipeSplitSink {
  const RowSink<T>& sink;
  const T* s_y;
  int32_t ra, nfin;  // nfin = rows that END in this tile; local row nfin (if any) is cut
  bool head_open;
  T* carry_out;
  T* head_val;
  __device__ __forceinline__ void put(int i, T sum) const {
    if (i == nfin) *carry_out = sum;
    else if (i == 0 && head_open) *head_val = sum;
    else sink.put_y(ra + i, sum, s_y[i]);
  }
};

template <typename T>
struct SplitArgs {  // only read by the SPLIT instantiation
  int32_t T_items, rows;
  int64_t nnz;
  T* carry_out;
  T* head_val;
  int32_t* head_row;
};

// CONSEC: CTA c walks a run of CONSECUTIVE tiles instead of c, c + grid, ... (see the host side)
template <typename T, int BLOCK, int NSTAGE, int G, bool SPLIT, bool CONSEC = false, bool FLAT = false>
__global__ void __launch_bounds__(BLOCK)
spmv_pipe_kernel(const int32_t* __restrict__ tile_row, const int32_t* __restrict__ tile_nnz,
                 int32_t n_tiles, const int32_t* __restrict__ rowptr,
                 const int32_t* __restrict__ colidx, const T* __restrict__ val,
                 const T* __restrict__ x, RowSink<T> sink, int32_t cap_nnz, int32_t cap_rows,
                 SplitArgs<T> sp) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ T red[2 * (BLOCK / 32)];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);     // [NSTAGE <= 8]
  int4* descs = reinterpret_cast<int4*>(smem_raw + 64);        // [NSTAGE] tile in each stage
  unsigned char* stages = smem_raw + kPipeHeaderBytes;
  const int tid = threadIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) mbar_init(&bars[s], BLOCK + 1);
    mbar_fence_init();
  }
  __syncthreads();
  int my_tiles;
  int64_t tile_base;
  if (CONSEC) {
    const int q_tiles = (int)n_tiles / (int)gridDim.x, r_tiles = (int)n_tiles % (int)gridDim.x;
    my_tiles = q_tiles + ((int)blockIdx.x < r_tiles ? 1 : 0);
    tile_base = (int64_t)blockIdx.x * q_tiles + min((int)blockIdx.x, r_tiles);
  } else {
    my_tiles = ((int)n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    tile_base = (int64_t)blockIdx.x;
  }
  auto tile_of = [&](int i) -> int64_t { return CONSEC ? tile_base + i : tile_base + (int64_t)i * gridDim.x; };

  // descriptor of my i-th tile: (r0, r1, n0, n1) = local rows [r0, r1), nnz window [n0, n1)
  auto load_desc = [&](int i) -> int4 {
    if (i >= my_tiles) return make_int4(0, 0, 0, 0);
    const int64_t b = tile_of(i);
    if (SPLIT) {
      // tile_row = merge-path split table (split_search_kernel): rows ra..rb-1 END in this tile,
      // row rb (if < rows) is cut by its end; nnz window from the merge diagonal
      const int32_t ra = __ldg(tile_row + b), rb = __ldg(tile_row + b + 1);
      const int64_t da = b * sp.T_items;
      const int64_t db = min(da + sp.T_items, (int64_t)sp.rows + sp.nnz);
      const int cut = rb < sp.rows ? 1 : 0;  // .y = local row count | cut flag << 24 (SPLIT only)
      return make_int4(ra, (rb - ra + cut) | (cut << 24), (int32_t)(da - ra), (int32_t)(db - rb));
    }
    return make_int4(__ldg(tile_row + b), __ldg(tile_row + b + 1), __ldg(tile_nnz + b),
                     __ldg(tile_nnz + b + 1));
  };
  // start every load of tile i (descriptor d) into stage i % NSTAGE
  auto issue = [&](int i, const int4& d) {
    if (i >= my_tiles) return;
    const int s = i % NSTAGE;
    uint64_t* bar = &bars[s];
    const PipeStage<T> st = pipe_stage<T>(stages, s, cap_nnz, cap_rows);
    const int32_t r0 = d.x, nrows = SPLIT ? (d.y & 0xffffff) : d.y - d.x, n0 = d.z, n1 = d.w;
    const int32_t n0a = n0 & ~3;
    const int32_t m0 = min((n0 + 3) & ~3, n1);  // aligned middle [m0, m1)
    const int32_t m1 = max(n1 & ~3, m0);
    if (tid == 0) {
      const uint32_t cnt = (uint32_t)(m1 - m0);
      descs[s] = d;  // published to the CTA by the release/acquire of the mbarrier phase
      mbar_arrive_expect_tx(bar, cnt * (uint32_t)(sizeof(T) + 4));
      if (cnt) {
        bulk_g2s(st.col + (m0 - n0a), colidx + m0, cnt * 4u, bar);
        bulk_g2s(st.val + (m0 - n0a), val + m0, cnt * (uint32_t)sizeof(T), bar);
      }
    } else if (tid >= 32 && tid < 32 + (m0 - n0)) {  // head, <= 3 elements
      const int32_t j = n0 + (tid - 32);
      cp_async4(st.col + (j - n0a), colidx + j);
      cp_async_elem(st.val + (j - n0a), val + j);
    } else if (tid >= 64 && tid < 64 + (n1 - m1)) {  // tail, <= 3 elements
      const int32_t j = m1 + (tid - 64);
      cp_async4(st.col + (j - n0a), colidx + j);
      cp_async_elem(st.val + (j - n0a), val + j);
    }
    for (int i2 = tid; i2 <= nrows; i2 += BLOCK) cp_async4(st.rp + i2, rowptr + r0 + i2);
    if (sink.stage_src)
      for (int i2 = tid; i2 < nrows; i2 += BLOCK) cp_async_elem(st.y + i2, sink.stage_src + r0 + i2);
    cp_async_mbar_arrive_noinc(bar);
  };

  int4 d_next = load_desc(0);
#pragma unroll 1
  for (int i = 0; i < NSTAGE - 1; ++i) {  // prologue: tiles 0 .. NSTAGE-2
    const int4 d = d_next;
    d_next = load_desc(i + 1);
    issue(i, d);
  }
  // d_next is now the descriptor of tile NSTAGE-1
#pragma unroll 1
  for (int i = 0; i < my_tiles; ++i) {
    {
      const int4 d = d_next;
      d_next = load_desc(i + NSTAGE);
      issue(i + NSTAGE - 1, d);
    }
    const int s = i % NSTAGE;
    const PipeStage<T> st = pipe_stage<T>(stages, s, cap_nnz, cap_rows);
    mbar_wait(&bars[s], (uint32_t)(i / NSTAGE) & 1u);
    const int4 d = descs[s];
    const int32_t r0 = d.x, nrows = SPLIT ? (d.y & 0xffffff) : d.y - d.x, n0 = d.z, n1 = d.w;
    if (SPLIT) {
      const int64_t b = tile_of(i);
      const bool has_cut = (d.y >> 24) != 0;
      const int32_t nfin = nrows - (has_cut ? 1 : 0);               // rows r0 .. r0+nfin-1 end here
      const bool head_open = nfin > 0 && st.rp[0] < n0;
      if (tid == 0) {
        sp.head_row[b] = head_open ? r0 : -1;
        if (!has_cut) sp.carry_out[b] = T(0);
      }
      if (nrows > 0) {
        // skewed rows: balance the x gathers first (every thread takes the same number of nnz,
        // products overwrite the staged values), then reduce the rows from shared memory
        const int32_t n0a = n0 & ~3;
        for (int j = (n0 - n0a) + tid; j < n1 - n0a; j += BLOCK * 4) {
          T vv[4], xx[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int jj = j + u * BLOCK;
            if (jj < n1 - n0a) { vv[u] = st.val[jj]; xx[u] = __ldg(x + st.col[jj]); }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (j + u * BLOCK < n1 - n0a) st.val[j + u * BLOCK] = mul_rn(vv[u], xx[u]);
        }
        __syncthreads();
        const PipeSplitSink<T> ps{sink, st.y, r0, has_cut ? nfin : -1, head_open, sp.carry_out + b, sp.head_val + b};
        reduce_rows_clamped_dispatch<T, BLOCK>(lanes_for_avg((n1 - n0) / nrows), st.val, st.rp, nrows, n0a, n0, n1, ps);
      }
    } else if (nrows > 0) {
      // rows straight from the staged tile: G lanes per row, U gathers in flight per lane
      const PipeSink<T> ps{sink, st.y, r0};
      if (FLAT)
        consume_rows_flat<T, G, BLOCK>(st.col, st.val, st.rp, nrows, n0, n1, x, ps);
      else
        consume_rows<T, G, (G == 1 ? 8 : 4), BLOCK, false>(st.col, st.val, st.rp, nrows, n0 & ~3, 0, 0, x, ps);
    }
    fence_proxy_async_smem();  // my generic-proxy accesses of this stage precede its next bulk fill
    __syncthreads();
  }
}

// ---- SELL: one pass over a column panel kept in sliced-ELL form (colpanel.cu) ---------------------------
// One warp per slice of 32 sorted rows, one CTA per window of 256 rows (8 slices).  Lane l walks ITS row front
// to back — entry u at sptr[slice] + u*32 + l, so the col_idx / val loads of a warp are single coalesced lines
// (streamed past L1), U of them and U x gathers in flight per lane — and adds the products in the row's
// storage order with separately rounded multiply and add, continuing from the sum carried over from the
// previous panels: bit-identical to spmv.h:24-30 when the row's columns ascend.  No shared memory, no
// barrier; the 256 KB of the SM stay L1 for the gathers in flight.  Padding entries carry col = -1.
// HINT: L2 eviction policies on the loads — the panel's col / val evict_first, the x gathers evict_last
template <typename T, int U, bool HINT = false>
__global__ void __launch_bounds__(kSellSigma)
spmv_sell_kernel(int32_t n_slices, const int32_t* __restrict__ sptr, const int32_t* __restrict__ perm,
                 const int32_t* __restrict__ col, const T* __restrict__ val, const T* __restrict__ x,
                 RowSink<T> sink, T* __restrict__ piece) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t pol_a = 0, pol_x = 0;
  if (HINT) { pol_a = l2_policy_evict_first(); pol_x = l2_policy_evict_last(); }
  constexpr int SPW = kSellSigma / kSellC;  // slices per window
  for (int64_t sl = (int64_t)blockIdx.x * SPW + warp; sl < n_slices; sl += (int64_t)gridDim.x * SPW) {
    const int32_t p0 = __ldg(sptr + sl), p1 = __ldg(sptr + sl + 1);
    const int32_t r = __ldg(perm + sl * kSellC + lane);
    T sum = (sink.phase >= 2 && r >= 0) ? sink.part[r] : T(0);
    for (int32_t p = p0 + lane; p < p1; p += U * kSellC) {
      int32_t c[U];
      T v[U], xv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        c[u] = -1;
        if (p + u * kSellC < p1) {
          c[u] = HINT ? ldg_stream_hint(col + p + u * kSellC, pol_a) : ldg_stream(col + p + u * kSellC);
          v[u] = HINT ? ldg_stream_hint(val + p + u * kSellC, pol_a) : ldg_stream(val + p + u * kSellC);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (c[u] >= 0) xv[u] = HINT ? ldg_hint(x + c[u], pol_x) : __ldg(x + c[u]);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (c[u] >= 0) sum = add_rn(sum, mul_rn(v[u], xv[u]));
    }
    if (r >= 0) {
      if (sink.phase == 0) sink.put(r, sum);   // a single panel: plain fused epilogue (reads y[r] itself)
      else sink.put_y(r, sum, T(0));           // phases 1-3 never use the staged value
    } else if (r <= -2) {
      piece[-(r + 2)] = sum;                   // virtual-row form: one piece of a cut row, finished by the fix-up kernel
    }
  }
}

// virtual-row form: adds the pieces of every cut row in a fixed order and finishes the row through the fused
// epilogue.  One THREAD per cut row while it has at most 32 pieces (most cut rows have 2-3: a warp per row spent its
// time in shuffle trees, 32 us on config 4); the few hub rows left are then taken one after the other by the whole
// warp (lane-strided partial sums + shuffle tree).
template <typename T>
__global__ void __launch_bounds__(256)
spmv_vsell_fixup_kernel(const int32_t* __restrict__ split, int32_t n_split, const T* __restrict__ piece, RowSink<T> sink) {
  const int lane = threadIdx.x & 31;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const bool valid = i < n_split;
  int32_t row = 0, p0 = 0, n = 0;
  if (valid) { row = split[3 * i]; p0 = split[3 * i + 1]; n = split[3 * i + 2]; }
  // phase 0: the fused epilogue; column-panel passes (phases 1-3): the pieces continue the sum carried in part[]
  auto emit = [&](int32_t r, T s) {
    if (sink.phase == 0) sink.put(r, s);
    else sink.put_y(r, sink.phase >= 2 ? add_rn(sink.part[r], s) : s, T(0));
  };
  if (valid && n <= 32) {
    T s = 0;
    for (int32_t k = 0; k < n; ++k) s += piece[p0 + k];
    emit(row, s);
  }
  unsigned m = __ballot_sync(0xffffffffu, valid && n > 32);
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const int32_t r2 = __shfl_sync(0xffffffffu, row, src), pp = __shfl_sync(0xffffffffu, p0, src), nn = __shfl_sync(0xffffffffu, n, src);
    T s = 0;
    for (int32_t k = lane; k < nn; k += 32) s += piece[pp + k];
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) s += __shfl_xor_sync(0xffffffffu, s, w);
    if (lane == 0) emit(r2, s);
  }
}

// G lanes per cut row (lane-strided partial sums, shuffle tree: a fixed order).  Measured on config 4 (same box, ms per
// step): 8 lanes per row 0.543, one warp per row 0.560, one thread per row 0.568 (profiles/r02_exp_c4_fixup.log)
template <typename T, int G>
__global__ void __launch_bounds__(256)
spmv_vsell_fixup_group_kernel(const int32_t* __restrict__ split, int32_t n_split, const T* __restrict__ piece, RowSink<T> sink) {
  const int lane = threadIdx.x & (G - 1);
  const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
  const bool valid = w < n_split;
  const int32_t row = valid ? split[3 * w] : 0, p0 = valid ? split[3 * w + 1] : 0, n = valid ? split[3 * w + 2] : 0;
  T s = 0;
  for (int32_t i = lane; i < n; i += G) s += piece[p0 + i];
#pragma unroll
  for (int m = G / 2; m > 0; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
  if (lane == 0 && valid) {
    if (sink.phase == 0) sink.put(row, s);
    else sink.put_y(row, sink.phase >= 2 ? add_rn(sink.part[row], s) : s, T(0));
  }
}

// rows [a, b) of the finished local y -> the same rows of every replica (peer mappings or one multicast address):
// consecutive lanes, consecutive elements — 128 / 256 contiguous bytes per warp store
template <typename T>
__global__ void __launch_bounds__(256)
spmv_replicate_kernel(const T* __restrict__ y, int32_t a, int32_t b, OutSpec<T> out) {
  for (int64_t i = a + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < b; i += (int64_t)gridDim.x * blockDim.x) {
    const T v = y[i];
    for (int d = 0; d < out.n; ++d) out.ptr[d][i] = v;
  }
}

// ---- VECTOR: G lanes per row, straight from global memory ----------------------------------
template <typename T, int G, int BLOCK, bool FMA>
__global__ void __launch_bounds__(BLOCK)
spmv_vector_kernel(int32_t rows, const int32_t* __restrict__ rowptr,
                   const int32_t* __restrict__ colidx, const T* __restrict__ val,
                   const T* __restrict__ x, RowSink<T> sink) {
  const int gl = threadIdx.x & (G - 1), gi = threadIdx.x / G;
  constexpr int RPB = BLOCK / G;
  for (int64_t rb = (int64_t)blockIdx.x * RPB; rb < rows; rb += (int64_t)gridDim.x * RPB) {
    const int64_t r = rb + gi;
    const bool valid = r < rows;
    const int s = valid ? rowptr[r] : 0, e = valid ? rowptr[r + 1] : 0;
    T sum = 0;
    for (int j = s + gl; j < e; j += G) {
      const T a = ldg_stream(val + j), b = __ldg(x + ldg_stream(colidx + j));
      sum = FMA ? fma_rn(a, b, sum) : add_rn(sum, mul_rn(a, b));
    }
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    if (valid && gl == 0) sink.put((int32_t)r, sum);
  }
}

// ---- SPLIT: merge-path tiles over (row ends + nnz); long rows are cut ------------------------
// split_row[b] = number of row ends consumed before diagonal b*T (Merrill & Garland's
// merge-path search); the nnz coordinate is b*T - split_row[b].
__global__ void split_search_kernel(const int32_t* __restrict__ rowptr, int32_t rows, int64_t nnz,
                                    int32_t T, int32_t n_tiles, int32_t* __restrict__ split_row) {
  int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > n_tiles) return;
  int64_t d = (int64_t)b * T;
  if (d > rows + nnz) d = rows + nnz;
  int64_t lo = d > nnz ? d - nnz : 0, hi = d < rows ? d : rows;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if ((int64_t)rowptr[mid + 1] <= d - mid - 1) lo = mid + 1; else hi = mid;
  }
  split_row[b] = (int32_t)lo;
}

template <typename T>
struct SplitSink {
  const RowSink<T>& sink;
  int32_t ra, nfin;
  bool head_open;  // row ra began in an earlier tile
  T* carry_out;    // this tile's slot
  T* head_val;
  __device__ __forceinline__ void put(int i, T sum) const {
    if (i == nfin) *carry_out = sum;                 // unfinished row rb -> next tiles
    else if (i == 0 && head_open) *head_val = sum;   // finished here, begun earlier -> fix-up
    else sink.put(ra + i, sum);
  }
};

// OVL: the tile's nnz window [ja, jb) follows from the split table alone, so the row pointers are
// fetched by LDGSTS (raw, clamped when read) and the old y rows are prefetched into L2 WHILE the A
// stream and the x gathers are in flight — two dependent DRAM round trips instead of four per CTA.
template <typename T, int BLOCK, bool VEC, bool OVL, bool NOL1 = false>
__global__ void __launch_bounds__(BLOCK)
spmv_split_kernel(const int32_t* __restrict__ split_row, int32_t T_items, int32_t rows, int64_t nnz,
                  const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                  const T* __restrict__ val, const T* __restrict__ x, RowSink<T> sink,
                  T* carry_out, T* head_val, int32_t* head_row, int32_t cap_nnz) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* s_prod = reinterpret_cast<T*>(smem_raw);
  int32_t* s_rp = reinterpret_cast<int32_t*>(s_prod + cap_nnz);
  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const int32_t ra = split_row[b], rb = split_row[b + 1];
  int64_t da = (int64_t)b * T_items, db = da + T_items;
  if (db > rows + nnz) db = rows + nnz;
  const int32_t ja = (int32_t)(da - ra), jb = (int32_t)(db - rb);
  const int32_t nfin = rb - ra;  // rows whose end lies in this tile
  const int nloc = rb < rows ? nfin + 1 : nfin;  // + the cut row, if there is one
  bool head_open;
  if (OVL) {
    for (int i = tid; i <= nfin; i += BLOCK) cp_async4(&s_rp[i], &rowptr[ra + i]);
    cp_async_commit();
    if (tid == 0) s_rp[nfin + 1] = jb;  // closes the cut row rb
    constexpr int kPer = 32 / (int)sizeof(T);  // one prefetch per 32-byte sector of y
    for (int i = tid * kPer; i < nfin; i += BLOCK * kPer) prefetch_l2(sink.y + ra + i);
    stage_products<T, BLOCK, VEC, NOL1>(colidx, val, x, ja, jb, s_prod);
    cp_async_wait<0>();
    __syncthreads();
    head_open = nfin > 0 && s_rp[0] < ja;
  } else {
    // local row pointers clamped to the tile's nnz window; entry nfin+1 closes the cut row rb
    for (int i = tid; i <= nfin; i += BLOCK) {
      int32_t v = rowptr[ra + i];
      s_rp[i] = v < ja ? ja : (v > jb ? jb : v);
    }
    if (tid == 0) s_rp[nfin + 1] = jb;
    __syncthreads();
    head_open = nfin > 0 && rowptr[ra] < ja;
    stage_products<T, BLOCK, VEC>(colidx, val, x, ja, jb, s_prod);
    __syncthreads();
  }
  if (tid == 0) head_row[b] = head_open ? ra : -1;
  if (rb >= rows && tid == 0) carry_out[b] = T(0);
  if (nloc <= 0) return;
  const SplitSink<T> ss{sink, ra, rb < rows ? nfin : -1, head_open, carry_out + b, head_val + b};
  if (OVL)
    reduce_rows_clamped_dispatch<T, BLOCK>(lanes_for_avg((jb - ja) / nloc), s_prod, s_rp, nloc, ja & ~3, ja, jb, ss);
  else
    reduce_rows_dispatch<T, BLOCK>(lanes_for_avg((jb - ja) / nloc), s_prod, s_rp, nloc, ja & ~3, ss);
}

// SPLIT, balanced flavour (SBLAS_FLAG_SPLIT_MERGE): same merge-path tiles, same carry / head tables and fix-up
// kernel, but phase 2 is ALSO a merge-path walk — thread t owns the merge items [t*I, (t+1)*I) of the
// tile (I = ceil(T_items / BLOCK) row ends or products, found by a binary search over the tile's
// clamped row pointers in shared memory) and sums its products in CSR order.  Rows that lie inside
// one thread are emitted directly; a row's last piece takes the partial sums of the threads before
// it from a CTA-wide SEGMENTED scan of the per-thread trailing sums (warp shuffles, then the eight
// warp totals), so a tile that is one slice of a 150 000-nnz row costs the same as a tile of 1 500
// short rows (the default phase 2 gives such a tile to ONE sub-warp while seven warps sit at the
// barrier).  Measured on config 4 (profiles/r01_sweep_c4_merge.log): 0.73 ms against 0.66 ms for the
// default at tile 1536 — the imbalance was not the bound, the divergent x gathers are — so opt-in only.
// Summation order: CSR order inside a thread, tree order across threads (deterministic).
template <typename T, int BLOCK, bool VEC>
__global__ void __launch_bounds__(BLOCK)
spmv_merge_kernel(const int32_t* __restrict__ split_row, int32_t T_items, int32_t rows, int64_t nnz,
                  const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                  const T* __restrict__ val, const T* __restrict__ x, RowSink<T> sink,
                  T* carry_out, T* head_val, int32_t* head_row, int32_t cap_nnz) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ T s_wv[BLOCK / 32];
  __shared__ int s_wf[BLOCK / 32];
  T* s_prod = reinterpret_cast<T*>(smem_raw);
  int32_t* s_rp = reinterpret_cast<int32_t*>(s_prod + cap_nnz);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int32_t ra = split_row[b], rb = split_row[b + 1];
  int64_t da = (int64_t)b * T_items, db = da + T_items;
  if (db > rows + nnz) db = rows + nnz;
  const int32_t ja = (int32_t)(da - ra), jb = (int32_t)(db - rb);
  const int32_t nfin = rb - ra;  // rows whose end lies in this tile
  // s_rp[i + 1] - ja = local nnz count at which row ra + i ends (i < nfin); clamped like the old kernel
  for (int i = tid; i <= nfin; i += BLOCK) {
    int32_t v = rowptr[ra + i];
    s_rp[i] = v < ja ? ja : (v > jb ? jb : v);
  }
  __syncthreads();
  const bool head_open = nfin > 0 && rowptr[ra] < ja;
  if (tid == 0) head_row[b] = head_open ? ra : -1;
  stage_products<T, BLOCK, VEC>(colidx, val, x, ja, jb, s_prod);
  __syncthreads();

  const int32_t Nl = jb - ja, total = nfin + Nl;
  const int32_t off = ja - (ja & ~3);  // products are parked at s_prod[j - (ja & ~3)]
  const int32_t items = (T_items + BLOCK - 1) / BLOCK;
  const int32_t d0 = min(tid * items, total), d1 = min(d0 + items, total);
  int lo = max(0, d0 - Nl), hi = min(d0, nfin);
  while (lo < hi) {  // row end `mid` comes before diagonal d0 iff all its nnz do (split_search_kernel's rule)
    const int mid = (lo + hi) >> 1;
    if (s_rp[mid + 1] - ja <= d0 - mid - 1) lo = mid + 1; else hi = mid;
  }
  int i = lo, j = d0 - lo;
  int e = i < nfin ? s_rp[i + 1] - ja : 0x7fffffff;
  T sum = 0, first_val = 0;
  int first_row = -1;
  for (int d = d0; d < d1; ++d) {
    if (j >= e) {  // row end: rows begun in this thread go out at once, the first one waits for its carry-in
      if (first_row < 0) { first_row = i; first_val = sum; }
      else sink.put(ra + i, sum);
      sum = 0;
      ++i;
      e = i < nfin ? s_rp[i + 1] - ja : 0x7fffffff;
    } else {
      sum = add_rn(sum, s_prod[off + j]);
      ++j;
    }
  }
  // segmented inclusive scan of (has_row_end, trailing sum): (f1,v1) (+) (f2,v2) = (f1|f2, f2 ? v2 : v1+v2)
  int f = first_row >= 0;
  T v = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const T pv = __shfl_up_sync(0xffffffffu, v, o);
    const int pf = __shfl_up_sync(0xffffffffu, f, o);
    if (lane >= o) {
      if (!f) v = add_rn(pv, v);
      f |= pf;
    }
  }
  T ev = __shfl_up_sync(0xffffffffu, v, 1);
  int ef = __shfl_up_sync(0xffffffffu, f, 1);
  if (lane == 0) { ev = 0; ef = 0; }
  if (lane == 31) { s_wv[warp] = v; s_wf[warp] = f; }
  __syncthreads();
  T pv = 0;
  for (int w = 0; w < warp; ++w) pv = s_wf[w] ? s_wv[w] : add_rn(pv, s_wv[w]);
  if (first_row >= 0) {
    const T carry_in = ef ? ev : add_rn(pv, ev);
    const T res = add_rn(carry_in, first_val);
    if (first_row == 0 && head_open) head_val[b] = res;  // begun in an earlier tile -> fix-up kernel
    else sink.put(ra + first_row, res);
  }
  if (tid == BLOCK - 1)  // what is left after the tile's last row end belongs to the cut row rb
    carry_out[b] = rb < rows ? (f ? v : add_rn(pv, v)) : T(0);
}

// SPLIT, segmented-sum flavour: phase 2 without a search and without global accesses in its loop.
//   (a) every non-empty row of the tile sets one bit (its first product) in a shared bitmap;
//   (b) thread t walks its I = ceil(T_items / BLOCK) consecutive products once, closing a segment
//       at every set bit: finished segments are written IN PLACE over the segment's first
//       product, the products before its first bit (`lead`) belong to a head in an earlier
//       thread, the last segment stays open;
//   (c) a reverse segmented scan over the threads' (has_bit, lead) pairs (warp shuffles + the eight
//       warp totals) hands every open segment the leads of the threads after it;
//   (d) one coalesced pass over the rows reads each row's sum from its first product slot and does
//       the fused y update.
// Balanced for any row-length mix (a tile that is one slice of a 150 000-nnz row costs the same as a
// tile of short rows), ~100 instructions per thread against ~250 for the sub-warp-per-row phase 2.
template <typename T, int BLOCK, bool VEC>
__global__ void __launch_bounds__(BLOCK)
spmv_split_seg_kernel(const int32_t* __restrict__ split_row, int32_t T_items, int32_t rows, int64_t nnz,
                      const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                      const T* __restrict__ val, const T* __restrict__ x, RowSink<T> sink,
                      T* carry_out, T* head_val, int32_t* head_row, int32_t cap_nnz) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ T s_wv[BLOCK / 32];
  __shared__ int s_wf[BLOCK / 32];
  T* s_prod = reinterpret_cast<T*>(smem_raw);
  int32_t* s_rp = reinterpret_cast<int32_t*>(s_prod + cap_nnz);
  uint32_t* s_flag = reinterpret_cast<uint32_t*>(s_rp + T_items + 3);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int32_t ra = split_row[b], rb = split_row[b + 1];
  int64_t da = (int64_t)b * T_items, db = da + T_items;
  if (db > rows + nnz) db = rows + nnz;
  const int32_t ja = (int32_t)(da - ra), jb = (int32_t)(db - rb);
  const int32_t nfin = rb - ra;
  const int nloc = rb < rows ? nfin + 1 : nfin;
  for (int i = tid; i <= nfin; i += BLOCK) cp_async4(&s_rp[i], &rowptr[ra + i]);
  cp_async_commit();
  if (tid == 0) s_rp[nfin + 1] = jb;
  for (int i = tid; i < (T_items + 31) / 32 + 1; i += BLOCK) s_flag[i] = 0u;
  constexpr int kPer = 32 / (int)sizeof(T);
  for (int i = tid * kPer; i < nfin; i += BLOCK * kPer) prefetch_l2(sink.y + ra + i);
  stage_products<T, BLOCK, VEC>(colidx, val, x, ja, jb, s_prod);
  cp_async_wait<0>();
  __syncthreads();
  const bool head_open = nfin > 0 && s_rp[0] < ja;
  if (tid == 0) head_row[b] = head_open ? ra : -1;
  if (rb >= rows && tid == 0) carry_out[b] = T(0);
  if (nloc <= 0) return;  // block-uniform
  // (a) one bit per non-empty row, at its first product (row pointers clamped to the tile's window)
  for (int i = tid; i < nloc; i += BLOCK) {
    const int s = min(max(s_rp[i], ja), jb) - ja, e = min(max(s_rp[i + 1], ja), jb) - ja;
    if (e > s) atomicOr(&s_flag[s >> 5], 1u << (s & 31));
  }
  __syncthreads();
  // (b)
  const int32_t Nl = jb - ja, off = ja - (ja & ~3);
  const int32_t items = (T_items + BLOCK - 1) / BLOCK;
  const int p0 = min(tid * items, Nl), p1 = min(p0 + items, Nl);
  T lead = 0, cur = 0;
  int head = -1;
  for (int p = p0; p < p1; ++p) {
    const bool fl = (s_flag[p >> 5] >> (p & 31)) & 1u;
    const T pr = s_prod[off + p];
    if (fl) {
      if (head >= 0) s_prod[off + head] = cur; else lead = cur;
      head = p;
      cur = pr;
    } else {
      cur = add_rn(cur, pr);
    }
  }
  if (head < 0) lead = cur;
  // (c) R_t = lead_t + (has_bit_t ? 0 : R_{t+1}); the open segment of thread t takes R_{t+1}
  int f = head >= 0;
  T v = lead;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const T nv = __shfl_down_sync(0xffffffffu, v, o);
    const int nf = __shfl_down_sync(0xffffffffu, f, o);
    if (lane + o < 32) {
      if (!f) v = add_rn(v, nv);
      f |= nf;
    }
  }
  T nvx = __shfl_down_sync(0xffffffffu, v, 1);
  int nfx = __shfl_down_sync(0xffffffffu, f, 1);
  if (lane == 31) { nvx = 0; nfx = 0; }
  if (lane == 0) { s_wv[warp] = v; s_wf[warp] = f; }
  __syncthreads();
  if (head >= 0) {
    T X = 0;  // suffix over the warps after this one
    for (int w = BLOCK / 32 - 1; w > warp; --w) X = s_wf[w] ? s_wv[w] : add_rn(s_wv[w], X);
    const T rest = nfx ? nvx : add_rn(nvx, X);
    s_prod[off + head] = add_rn(cur, rest);
  }
  __syncthreads();
  // (d)
  const SplitSink<T> ss{sink, ra, rb < rows ? nfin : -1, head_open, carry_out + b, head_val + b};
  for (int i = tid; i < nloc; i += BLOCK) {
    const int s = min(max(s_rp[i], ja), jb) - ja, e = min(max(s_rp[i + 1], ja), jb) - ja;
    ss.put(i, e > s ? s_prod[off + s] : T(0));
  }
}

template <typename T>
__global__ void spmv_split_fixup_kernel(const int32_t* __restrict__ rowptr, int32_t T_items,
                                        int32_t n_tiles, const T* __restrict__ carry_out,
                                        const T* __restrict__ head_val,
                                        const int32_t* __restrict__ head_row, RowSink<T> sink) {
  int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_tiles) return;
  const int32_t hr = head_row[b];
  if (hr < 0) return;
  // first tile holding nnz of row hr: merge coordinate of its first nnz = hr + rowptr[hr]
  const int32_t bs = (int32_t)(((int64_t)hr + rowptr[hr]) / T_items);
  T sum = 0;
  for (int32_t t = bs; t < b; ++t) sum += carry_out[t];
  sum += head_val[b];
  sink.put(hr, sum);
}

// ---- host side ---------------------------------------------------------------------------------
template <typename T, int BLOCK, bool FMA>
static int launch_vector(cudaStream_t st, int lanes, int32_t rows, const int32_t* rp,
                         const int32_t* ci, const T* v, const T* x, const RowSink<T>& sink) {
  auto grid = [&](int G) { return (int)std::min<int64_t>(ceil_div(rows, BLOCK / G), kNumSM * 32); };
  switch (lanes) {
#define SBLAS_CASE(G)                                                                        \
  case G: spmv_vector_kernel<T, G, BLOCK, FMA><<<grid(G), BLOCK, 0, st>>>(rows, rp, ci, v, x, sink); break;
    SBLAS_CASE(1) SBLAS_CASE(2) SBLAS_CASE(4) SBLAS_CASE(8) SBLAS_CASE(16)
    default: spmv_vector_kernel<T, 32, BLOCK, FMA><<<grid(32), BLOCK, 0, st>>>(rows, rp, ci, v, x, sink);
#undef SBLAS_CASE
  }
  SBLAS_LAUNCH_CHECK();
  return 0;
}

static int pow2_lanes(double avg) {
  int l = 1;
  while (l < 32 && l * 4 < avg) l <<= 1;  // ~4+ nnz per lane
  return l;
}

static int pow2_lanes_pipe(double avg) {
  return avg <= 6 ? 1 : avg <= 12 ? 2 : avg <= 24 ? 4 : avg <= 48 ? 8 : avg <= 96 ? 16 : 32;
}

template <typename T, bool SPLIT, bool CONSEC = false>
static const void* pipe_kernel_ptr(int stages, int G) {
#define SBLAS_PK(NS, GG) (const void*)spmv_pipe_kernel<T, kSpmvBlock, NS, GG, SPLIT, CONSEC>
#define SBLAS_PKG(NS)                                                                          \
  (G <= 1 ? SBLAS_PK(NS, 1) : G == 2 ? SBLAS_PK(NS, 2) : G == 4 ? SBLAS_PK(NS, 4)               \
   : G == 8 ? SBLAS_PK(NS, 8) : G == 16 ? SBLAS_PK(NS, 16) : SBLAS_PK(NS, 32))
  return stages <= 2 ? SBLAS_PKG(2) : stages == 3 ? SBLAS_PKG(3) : SBLAS_PKG(4);
#undef SBLAS_PKG
#undef SBLAS_PK
}

// short-row flavour (consume_rows_flat): 2 stages, strided tile order; G = lanes per row of its reduction phase
template <typename T>
static const void* pipe_flat_kernel_ptr(int G) {
#define SBLAS_PX(GG) (const void*)spmv_pipe_kernel<T, kSpmvBlock, 2, GG, false, false, true>
  return G <= 2 ? SBLAS_PX(2) : G == 4 ? SBLAS_PX(4) : G == 8 ? SBLAS_PX(8) : G == 16 ? SBLAS_PX(16) : SBLAS_PX(32);
#undef SBLAS_PX
}

// Column locality of consecutive rows, sampled: sum over ~4096 evenly spaced rows r (r and r+1 both non-empty)
// of |colidx[rowptr[r+1]] - colidx[rowptr[r]]|, and the number of samples.
__global__ void col_jump_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                                int32_t rows, int32_t n_samples, double* __restrict__ out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  double jump = 0.0, cnt = 0.0;
  if (s < n_samples) {
    const int64_t r = (int64_t)s * (rows - 1) / n_samples;
    if (r + 1 < rows) {
      const int32_t a0 = rowptr[r], a1 = rowptr[r + 1], a2 = rowptr[r + 2];
      if (a1 > a0 && a2 > a1) {
        jump = fabs((double)colidx[a1] - (double)colidx[a0]);
        cnt = 1.0;
      }
    }
  }
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) {
    jump += __shfl_xor_sync(0xffffffffu, jump, m);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, m);
  }
  if ((threadIdx.x & 31) == 0 && cnt > 0.0) {
    atomicAdd(out, jump);
    atomicAdd(out + 1, cnt);
  }
}

constexpr size_t kGatherHeavyXBytes = 96u << 20;  // x beyond this does not stay in the 126 MB L2 beside the A stream

// Gather-heavy inputs: x does not fit L2 and consecutive rows look at unrelated parts of it (uniform-random
// columns), so every gather is a DRAM sector of its own.  sync2gpu answers with the column-panel mirror
// (colpanel.cu); without it AUTO takes the sub-warp-per-row VECTOR kernel, which keeps more gathers in
// flight than PIPE's staged tiles (16M x 32 uniform-random, fp64: 5.6 ms against 7.4 ms).  The statistic
// costs one tiny kernel + one 16-byte copy and a stream synchronisation, once per plan: sblas_plan_prepare_i32
// / sblas_plan_build_colpanels_* take it at sync2gpu time so that no SpMV call has to.
bool measure_gather_heavy(cudaStream_t st, sblasPlan_st* plan, const int32_t* rp, const int32_t* ci, int32_t rows,
                          int32_t cols, size_t elem_size) {
  if ((size_t)cols * elem_size <= kGatherHeavyXBytes || rows < 4096 || plan->nnz == 0) return false;
  if (plan->col_jump < 0.0) {
    void* scratch = nullptr;
    if (plan_scratch(plan, 64, &scratch)) return false;
    double h[2] = {0.0, 0.0};
    if (cudaMemsetAsync(scratch, 0, 16, st) != cudaSuccess) return false;
    col_jump_kernel<<<16, 256, 0, st>>>(rp, ci, rows, 4096, (double*)scratch);
    ++g_launches;
    if (cudaMemcpyAsync(h, scratch, 16, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess)
      return false;
    plan->col_jump = h[1] > 0.0 ? h[0] / h[1] : 0.0;
  }
  return plan->col_jump > (double)cols / 256.0;
}

template <typename T>
static int resolve_variant(const sblasPlan_st* plan, const sblasOpts* o, int32_t tile) {
  if (o && o->variant != SBLAS_VARIANT_AUTO) return o->variant;
  if (!plan) return SBLAS_SPMV_VECTOR;
  (void)tile;
  return plan->max_row_nnz <= kSpmvPipeMaxRow ? SBLAS_SPMV_PIPE : SBLAS_SPMV_SPLIT;
}

template <typename T>
static int set_smem_once(const void* fn, int bytes) {
  if (bytes > 48 * 1024) {
    SBLAS_CHECK(bytes <= 227 * 1024, "spmv: tile needs %d bytes of shared memory", bytes);
    SBLAS_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  }
  return 0;
}

// tile (merge items) PIPE uses for a plan: every row must fit a stage's staging buffer
static int32_t pipe_tile_for(const sblasPlan_st* plan) {
  if (plan->pipe_tile > 0) return plan->pipe_tile;
  return std::max<int32_t>(kSpmvPipeDefaultTile,
                           std::min<int32_t>(2 * kSpmvPipeMaxRow, (2 * plan->max_row_nnz + 511) & ~511));
}

int prepare_pipe_table(cudaStream_t st, sblasPlan_st* plan, bool* pipe_ok) {
  if (pipe_ok) *pipe_ok = plan->max_row_nnz <= kSpmvPipeMaxRow;
  if (plan->max_row_nnz > kSpmvPipeMaxRow) return 0;  // SPLIT builds its own table
  const TileTable* tt = nullptr;
  return plan_get_table(plan, st, 1, pipe_tile_for(plan), &tt);
}

// PIPE launch on one CSR (the whole local matrix, or one column panel of it).  *fallback = the variant to take
// instead when this input cannot be staged (unaligned col/val -> TILE, a row too long -> SPLIT).
template <typename T>
static int launch_pipe(cudaStream_t st, sblasPlan_st* plan, int32_t rows, int64_t nnz, const int32_t* rp,
                       const int32_t* ci, const T* v, const T* x, const RowSink<T>& sink, const sblasOpts* o,
                       bool auto_gather, int* fallback, bool short_rows = false) {
  *fallback = 0;
  if (!(aligned16(ci) && aligned16(v))) { *fallback = SBLAS_SPMV_TILE; return 0; }
  const int32_t tile = (o && o->tile > 0) ? o->tile : pipe_tile_for(plan);
  const TileTable* tt = nullptr;
  int rc = plan_get_table(plan, st, 1, tile, &tt);
  if (rc) return rc;
  const int32_t cap_nnz = (tt->max_tile_nnz + 8 + 3) & ~3;  // + head alignment slack
  const int32_t cap_rows = tt->max_tile_rows + 1;
  int stages = (o && o->stages > 0) ? o->stages : 2;
  stages = std::min(std::max(stages, 2), 4);
  auto smem_for = [&](int ns) { return (size_t)kPipeHeaderBytes + ns * pipe_stage_bytes<T>(cap_nnz, cap_rows); };
  while (stages > 2 && smem_for(stages) > 110 * 1024) --stages;  // keep >= 2 CTAs per SM
  if (smem_for(stages) > 200 * 1024) { *fallback = SBLAS_SPMV_SPLIT; return 0; }  // a row too long: cut rows instead
  const size_t smem = smem_for(stages);
  // lanes per row from the mean row length (one instantiation per launch keeps the
  // kernel small: ~50 registers instead of 160 with an in-kernel switch)
  int G = (o && o->lanes > 0) ? o->lanes : pow2_lanes_pipe(rows ? (double)(nnz / rows) : 0.0);  // integer mean, as the per-tile rule of TILE
  // column-panel passes (short_rows): the two-phase flat consumer, its reduction phase with one lane group per
  // row as wide as the mean row; opts->lanes = 1 keeps the one-lane-per-row kernel (the CPU loop's order)
  const bool flat = short_rows && !(o && o->lanes == 1) && stages == 2;
  if (flat && !(o && o->lanes > 1)) {
    const double mean = rows ? (double)nnz / rows : 0.0;
    G = 2;
    while (G < 32 && G < mean) G <<= 1;
  }
  // tile order: CTA c takes tiles c, c+grid, ... (the resident CTAs sweep one contiguous window of A and
  // share x through L2: best while x fits L2, c2 74 us against 86 us) or a run of consecutive tiles (the x
  // lines of neighbouring tiles stay in the CTA's own L1: best when x is larger than L2 and rows are local,
  // b32 1.376 -> 1.323 ms).  The part of x these rows touch is estimated as rows * col_jump elements (local
  // rows advance col_jump columns per row; a row panel of a larger matrix touches only its own window of x);
  // col_jump is only measured when the whole x exceeds 96 MB.  SBLAS_PIPE_CONSEC=0/1 overrides
  // for experiments.  A template parameter: the strided instantiation is the kernel it always was.
  static const int env_consec = [] { const char* e = getenv("SBLAS_PIPE_CONSEC"); return e ? atoi(e) : -1; }();
  const bool big_window = plan->col_jump >= 0.0 && !auto_gather &&
                          plan->col_jump * (double)rows * sizeof(T) > (double)kGatherHeavyXBytes;
  const bool consec = env_consec >= 0 ? env_consec != 0 : big_window;
  const void* kern = flat ? pipe_flat_kernel_ptr<T>(G)
                          : consec ? pipe_kernel_ptr<T, false, true>(stages, G) : pipe_kernel_ptr<T, false>(stages, G);
  int per_sm = 0, n_sm = kNumSM;
  if ((rc = kernel_residency(kern, kSpmvBlock, smem, &per_sm, &n_sm)) != 0) return rc;
  const int grid = std::min<int>(tt->n_tiles, n_sm * per_sm);
  SplitArgs<T> sp{};
  void* args[] = {(void*)&tt->d_tile_row, (void*)&tt->d_tile_nnz, (void*)&tt->n_tiles, (void*)&rp,
                  (void*)&ci, (void*)&v, (void*)&x, (void*)&sink, (void*)&cap_nnz,
                  (void*)&cap_rows, (void*)&sp};
  cudaError_t e = cudaLaunchKernel(kern, dim3(grid), dim3(kSpmvBlock), args, smem, st);
  ++g_launches;
  if (e != cudaSuccess) { set_error("spmv pipe launch: %s", cudaGetErrorString(e)); return 3; }
  return 0;
}

template <typename T>
static int spmv_impl(cudaStream_t st, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                     const int32_t* ci, const T* v, const T* x, T* y, T alpha, T beta,
                     sblasPlan_st* plan, const sblasOpts* o) {
  SBLAS_CHECK(rows >= 0 && cols >= 0 && nnz >= 0 && nnz < (1ll << 31), "spmv: bad sizes");
  if (rows == 0) return 0;
  SBLAS_CHECK(rp && y && (nnz == 0 || (ci && v && x)), "spmv: null pointer");
  const int flags = o ? o->flags : 0;
  SBLAS_CHECK(!(flags & (SBLAS_FLAG_HEAD_PARTIAL | SBLAS_FLAG_TAIL_PARTIAL)) || (o && o->carry),
              "spmv: HEAD/TAIL_PARTIAL needs opts->carry");
  RowSink<T> sink;
  sink.y = y;
  sink.out = make_outspec<T>(o);
  sink.carry = o ? (T*)o->carry : nullptr;
  sink.rows = rows;
  sink.flags = flags;
  sink.alpha = alpha;
  sink.beta = beta;
  sink.part = nullptr;
  sink.stage_src = y;
  sink.phase = 0;
  int32_t tile = (o && o->tile > 0) ? o->tile : kSpmvDefaultTile;
  EphemeralPlan eph;  // released on every exit path
  eph.st = st;
  int variant = resolve_variant<T>(plan, o, tile);
  if (!plan && variant != SBLAS_SPMV_VECTOR) {
    if (sblas_plan_create_i32(st, rows, nnz, rp, &eph.p)) return 4;
    plan = eph.p;
  }
  if (plan) SBLAS_CHECK(plan->rows == rows && plan->nnz == nnz, "spmv: plan does not match matrix");
  const bool user_variant = o && o->variant != SBLAS_VARIANT_AUTO;
  const bool have_panels = plan && plan->cp.n > 0 && plan->cp.is_f64 == (sizeof(T) == 8) &&
                           (int64_t)cols <= (int64_t)plan->cp.n * plan->cp.panel_cols;
  if (variant == SBLAS_SPMV_COLPANEL)
    SBLAS_CHECK(have_panels, "spmv: SBLAS_SPMV_COLPANEL needs a plan with column panels of this value type (sblas_plan_build_colpanels_*)");
  bool auto_gather = false;
  // a skewed input's virtual-row panel finishes rows in sorted order: as REMOTE stores (replicas) every 4-byte
  // element would be an NVLink transaction of its own, and a separate coalesced copy afterwards does not overlap
  // with the compute (config 4, 2 / 4 GPUs: SPLIT with in-kernel peer stores 0.364 / 0.203 ms, the mirror with
  // scattered stores 0.356 / 0.406 ms, with a copy kernel afterwards 0.406 / 0.272 ms): with replicas AUTO keeps SPLIT
  const bool skip_mirror = have_panels && plan->cp.skewed && sink.out.n > 0;
  if (!user_variant && have_panels && !skip_mirror) {
    variant = SBLAS_SPMV_COLPANEL;
  } else if (plan && variant == SBLAS_SPMV_PIPE && !user_variant &&
             measure_gather_heavy(st, plan, rp, ci, rows, cols, sizeof(T))) {
    variant = SBLAS_SPMV_VECTOR;
    auto_gather = true;
  }
  const bool vec = aligned16(ci) && aligned16(v);
  // SPLIT reads col_idx / val ONE nnz per lane and load (consecutive lanes = consecutive nnz, coalesced 4-byte loads)
  // rather than four consecutive nnz per lane by 128-bit loads: the 32 x gathers of one instruction then belong to
  // 32 CONSECUTIVE nnz — mostly one row, ascending columns — and share x lines where the row is dense (the hub
  // columns of a power-law matrix), while at stride 4 every gather instruction touches the lines of 128 nnz.  The
  // L1TEX wavefronts of the gathers are what bounds config 4: 0.632 -> 0.598 ms (profiles/r02_exp_c4_scalar.log).
  // SBLAS_SPLIT_VEC=1 brings the 128-bit loads back.
  static const int env_split_vec = [] { const char* e = getenv("SBLAS_SPLIT_VEC"); return e ? atoi(e) : 0; }();
  const bool vec_split = vec && env_split_vec != 0;
  int rc = 0;
  if (variant == SBLAS_SPMV_COLPANEL) {
    // P passes of the PIPE kernel, one per column panel of the plan's mirror (colpanel.cu): pass p gathers
    // only from x[p*W, (p+1)*W), which stays L2-resident while the panel streams by
    const ColPanels& cp = plan->cp;
    sblasOpts po;
    if (o) { po = *o; po.tile = 0; }   // the panels' tile tables were sized at build time
    // sliced-ELL passes finish rows in sorted order (windows of up to 65 536 rows) and store them to the replicas from
    // the kernel, overlapped with the compute.  SBLAS_SELL_REPLICATE_AFTER=1: the passes write the local y only and one
    // coalesced copy kernel pushes the finished rows to every replica afterwards (measured slower on config 4: the
    // copy does not overlap).
    static const int env_rep_after = [] { const char* e = getenv("SBLAS_SELL_REPLICATE_AFTER"); return e ? atoi(e) : 0; }();
    const bool replicate_after = env_rep_after != 0 && sink.out.n > 0 && !cp.sell.empty();   // opt-in (experiments)
    for (int p = 0; p < cp.n && rc == 0; ++p) {
      RowSink<T> ps = sink;
      if (replicate_after) ps.out.n = 0;
      ps.part = (T*)cp.d_part;
      ps.phase = cp.n == 1 ? 0 : p == 0 ? 1 : p == cp.n - 1 ? 3 : 2;
      ps.stage_src = ps.phase == 0 ? y : ps.phase == 1 ? nullptr : (const T*)cp.d_part;
      if (!cp.sell.empty()) {
        // sliced-ELL panels: one CTA per window of 256 rows, persistent over the windows
        const ColPanels::Sell& sl = cp.sell[(size_t)p];
        const int n_win = sl.n_slices / (kSellSigma / kSellC);
        // entries in flight per lane: SBLAS_SELL_U = 2 / 4 / 8 (default 8: c3v 2.95 / 2.83 / 2.77 ms, profiles/r02_exp_c3v_sell_u_grid.log)
        static const int env_u = [] { const char* e = getenv("SBLAS_SELL_U"); return e ? atoi(e) : 8; }();
        const void* kern = env_u == 8 ? (const void*)spmv_sell_kernel<T, 8> : env_u == 2 ? (const void*)spmv_sell_kernel<T, 2>
                                                                                          : (const void*)spmv_sell_kernel<T, 4>;
        int per_sm = 0, n_sm = kNumSM;
        if ((rc = kernel_residency(kern, kSellSigma, 0, &per_sm, &n_sm)) != 0) break;
        static const int env_g = [] { const char* e = getenv("SBLAS_SELL_GRID"); return e ? atoi(e) : 0; }();   // CTAs per SM (experiments)
        const int grid = std::max(1, std::min(n_win, n_sm * (env_g > 0 ? std::min(env_g, per_sm) : per_sm)));
        // (forcing 5 / 6 resident CTAs with __launch_bounds__ spills at 48 / 40 registers: 2.97 -> 3.73 / 4.07 ms, dropped)
        // multi-panel passes: L2 eviction policies on the loads (panel stream evict_first, x gathers evict_last) keep
        // the x slice resident while the panel streams by: c3v 2.69 -> 2.65 ms at 3 panels, 3.10 -> 3.00 at 2
        // (profiles/r02_exp_c3v_l2hint.log).  SBLAS_SELL_L2HINT=0 turns them off.
        static const int env_hint = [] { const char* e = getenv("SBLAS_SELL_L2HINT"); return e ? atoi(e) : 1; }();
        if (env_hint && cp.n > 1 && env_u == 8)
          spmv_sell_kernel<T, 8, true><<<grid, kSellSigma, 0, st>>>(sl.n_slices, sl.d_sptr, sl.d_perm, sl.d_col, (const T*)sl.d_val, x, ps, (T*)sl.d_piece);
        else
        if (env_u == 8) spmv_sell_kernel<T, 8><<<grid, kSellSigma, 0, st>>>(sl.n_slices, sl.d_sptr, sl.d_perm, sl.d_col, (const T*)sl.d_val, x, ps, (T*)sl.d_piece);
        else if (env_u == 2) spmv_sell_kernel<T, 2><<<grid, kSellSigma, 0, st>>>(sl.n_slices, sl.d_sptr, sl.d_perm, sl.d_col, (const T*)sl.d_val, x, ps, (T*)sl.d_piece);
        else
        spmv_sell_kernel<T, 4><<<grid, kSellSigma, 0, st>>>(sl.n_slices, sl.d_sptr, sl.d_perm, sl.d_col, (const T*)sl.d_val, x, ps, (T*)sl.d_piece);
        ++g_launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("spmv sell launch: %s", cudaGetErrorString(e)); rc = 3; }
        if (rc == 0 && sl.n_split > 0) {
          // lanes per cut row: SBLAS_VSELL_FIXUP = 8 (default) / 32 / 1 (a thread per row, the warp for hub rows)
          static const int fix_g = [] { const char* e = getenv("SBLAS_VSELL_FIXUP"); return e ? atoi(e) : 8; }();
          if (fix_g == 1)
            spmv_vsell_fixup_kernel<T><<<(int)ceil_div((int64_t)sl.n_split, 256), 256, 0, st>>>(sl.d_split, sl.n_split, (const T*)sl.d_piece, ps);
          else if (fix_g == 32)
            spmv_vsell_fixup_group_kernel<T, 32><<<(int)ceil_div((int64_t)sl.n_split * 32, 256), 256, 0, st>>>(sl.d_split, sl.n_split, (const T*)sl.d_piece, ps);
          else
            spmv_vsell_fixup_group_kernel<T, 8><<<(int)ceil_div((int64_t)sl.n_split * 8, 256), 256, 0, st>>>(sl.d_split, sl.n_split, (const T*)sl.d_piece, ps);
          ++g_launches;
          e = cudaGetLastError();
          if (e != cudaSuccess) { set_error("spmv vsell fixup launch: %s", cudaGetErrorString(e)); rc = 3; }
        }
        continue;
      }
      sblasPlan_st* sub = cp.sub[(size_t)p];
      const int64_t off = cp.nnz_off[(size_t)p];
      int fb = 0;
      rc = launch_pipe<T>(st, sub, rows, sub->nnz, sub->d_rowptr, cp.d_colidx + off, (const T*)cp.d_val + off, x, ps,
                          o ? &po : nullptr, false, &fb, true);
      if (rc == 0 && fb != 0) { set_error("spmv: column panel %d cannot be staged by the PIPE kernel", p); rc = 5; }
    }
    if (rc == 0 && replicate_after) {
      // rows that went to the carry buffer (partial first / last row of a segment) are not the replicas' business
      const int32_t a = (flags & SBLAS_FLAG_HEAD_PARTIAL) ? 1 : 0, b = rows - ((flags & SBLAS_FLAG_TAIL_PARTIAL) ? 1 : 0);
      if (b > a) {
        const int grid = (int)std::min<int64_t>(ceil_div((int64_t)(b - a), 256), (int64_t)kNumSM * 16);
        spmv_replicate_kernel<T><<<grid, 256, 0, st>>>(y, a, b, sink.out);
        ++g_launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("spmv replicate launch: %s", cudaGetErrorString(e)); rc = 3; }
      }
    }
    return rc;
  }
  if (variant == SBLAS_SPMV_PIPE) {
    // persistent TMA-staged tiles; needs 16-byte aligned col_idx / val (else: TILE)
    int fb = 0;
    rc = launch_pipe<T>(st, plan, rows, nnz, rp, ci, v, x, sink, o, auto_gather, &fb);
    if (rc == 0 && fb != 0) variant = fb;
  }
  if (rc == 0 && variant == SBLAS_SPMV_TILE) {
    const TileTable* tt = nullptr;
    if ((rc = plan_get_table(plan, st, 1, tile, &tt)) == 0) {
      const int32_t cap = (tt->max_tile_nnz + 9) & ~3;
      const int smem = cap * (int)sizeof(T) + (tt->max_tile_rows + 2) * (int)sizeof(int32_t);
      if (smem > kSpmvMaxSmem) {
        variant = SBLAS_SPMV_SPLIT;  // a row too long for one tile: cut rows instead
      } else {
        auto kern = vec ? spmv_tile_kernel<T, kSpmvBlock, true> : spmv_tile_kernel<T, kSpmvBlock, false>;
        if ((rc = set_smem_once<T>((const void*)kern, smem)) == 0) {
          kern<<<tt->n_tiles, kSpmvBlock, smem, st>>>(tt->d_tile_row, rp, ci, v, x, sink, cap,
                                                      o ? o->lanes : 0);
          ++g_launches;
          cudaError_t e = cudaGetLastError();
          if (e != cudaSuccess) { set_error("spmv tile launch: %s", cudaGetErrorString(e)); rc = 3; }
        }
      }
    }
  }
  if (rc == 0 && variant == SBLAS_SPMV_SPLIT) {
    if (!(o && o->tile > 0)) tile = kSpmvSplitDefaultTile;
    // table (row_weight = -1) marks the merge-path split table; built here, cached in the plan
    const TileTable* tt = nullptr;
    for (const TileTable& t : plan->tables)
      if (t.tile == tile && t.row_weight == -1) tt = &t;
    if (!tt) {
      TileTable t;
      t.tile = tile;
      t.row_weight = -1;
      t.d_tile_nnz = nullptr;  // the split table has no nnz column: windows follow from the merge diagonal
      t.n_tiles = (int32_t)ceil_div(std::max<int64_t>((int64_t)rows + nnz, 1), tile);
      SBLAS_CUDA(cudaMalloc(&t.d_tile_row, sizeof(int32_t) * (size_t)(t.n_tiles + 1)));
      split_search_kernel<<<(int)ceil_div(t.n_tiles + 1, 256), 256, 0, st>>>(rp, rows, nnz, tile,
                                                                           t.n_tiles, t.d_tile_row);
      ++g_launches;
      if (cudaGetLastError() != cudaSuccess) {
        cudaFree(t.d_tile_row);
        set_error("spmv split: table build launch failed");
        return 3;
      }
      t.max_tile_nnz = tile;
      t.max_tile_rows = tile;
      plan->tables.reserve(16);
      plan->tables.push_back(t);
      tt = &plan->tables.back();
    }
    void* scratch = nullptr;
    const size_t per = 2 * sizeof(T) + sizeof(int32_t);
    if ((rc = plan_scratch(plan, per * (size_t)tt->n_tiles + 64, &scratch)) == 0) {
      T* carry_out = (T*)scratch;
      T* head_val = carry_out + tt->n_tiles;
      int32_t* head_row = (int32_t*)(head_val + tt->n_tiles);
      if (vec && o && o->stages > 0) {
        // staged (PIPE) flavour, on request (opts->stages): same tiles, TMA bulk staging, persistent
        // CTAs.  Measured slower than the one-tile-per-CTA kernel below on power-law inputs
        // (config 4: 0.94 ms against 0.66 ms), so AUTO does not take it.
        const int32_t cap_nnz = (tile + 8 + 3) & ~3;
        const int32_t cap_rows = tile + 2;
        int stages = (o && o->stages > 0) ? std::min(std::max(o->stages, 2), 4) : 2;
        auto smem_for = [&](int ns) { return (size_t)kPipeHeaderBytes + ns * pipe_stage_bytes<T>(cap_nnz, cap_rows); };
        while (stages > 2 && smem_for(stages) > 110 * 1024) --stages;
        const size_t smem = smem_for(stages);
        int G = (o && o->lanes > 0) ? o->lanes : pow2_lanes_pipe(rows ? (double)(nnz / std::max<int64_t>(rows - plan->empty_rows, 1)) : 0.0);
        const void* kern = pipe_kernel_ptr<T, true>(stages, G);
        int per_sm = 0, n_sm = kNumSM;
        if (smem > 200 * 1024) { set_error("spmv split: tile %d does not fit shared memory", tile); rc = 2; }
        if (rc == 0) rc = kernel_residency(kern, kSpmvBlock, smem, &per_sm, &n_sm);
        if (rc == 0) {
          const int grid = std::min<int>(tt->n_tiles, n_sm * per_sm);
          SplitArgs<T> sp{tile, rows, nnz, carry_out, head_val, head_row};
          const int32_t* no_nnz_table = nullptr;
          void* args[] = {(void*)&tt->d_tile_row, (void*)&no_nnz_table, (void*)&tt->n_tiles, (void*)&rp,
                          (void*)&ci, (void*)&v, (void*)&x, (void*)&sink, (void*)&cap_nnz,
                          (void*)&cap_rows, (void*)&sp};
          cudaError_t e = cudaLaunchKernel(kern, dim3(grid), dim3(kSpmvBlock), args, smem, st);
          ++g_launches;
          if (e != cudaSuccess) { set_error("spmv split(pipe) launch: %s", cudaGetErrorString(e)); rc = 3; }
        }
      } else {
        const int32_t cap = (tile + 9) & ~3;
        const int smem = cap * (int)sizeof(T) + (tile + 3) * (int)sizeof(int32_t) + ((tile + 31) / 32 + 2) * 4;
        // phase 2: sub-warp per row (default) or the balanced merge-path walk (opt-in: it removes the idle
        // warps on tiles that are one slice of a long row, but its search + scan cost more than they did)
        const bool legacy = (flags & SBLAS_FLAG_SPLIT_MERGE) == 0;
        const bool serial = (flags & SBLAS_FLAG_SPLIT_SERIAL) != 0;
        auto kern = legacy ? (serial ? (vec_split ? spmv_split_kernel<T, kSpmvBlock, true, false> : spmv_split_kernel<T, kSpmvBlock, false, false>)
                                     : (flags & SBLAS_FLAG_NO_L1)
                                           ? (vec_split ? spmv_split_kernel<T, kSpmvBlock, true, true, true> : spmv_split_kernel<T, kSpmvBlock, false, true, true>)
                                           : (vec_split ? spmv_split_kernel<T, kSpmvBlock, true, true> : spmv_split_kernel<T, kSpmvBlock, false, true>))
                           : (vec_split ? spmv_merge_kernel<T, kSpmvBlock, true> : spmv_merge_kernel<T, kSpmvBlock, false>);
        if (flags & SBLAS_FLAG_SPLIT_SEG)
          kern = vec_split ? spmv_split_seg_kernel<T, kSpmvBlock, true> : spmv_split_seg_kernel<T, kSpmvBlock, false>;
        {  // experiment hook: SBLAS_SPLIT_CARVEOUT=<percent of the 256 KB given to shared memory>; the driver's
           // default takes whatever makes the most CTAs resident, which can leave almost no L1 (tile >= 3072)
          static const int carve = [] { const char* e = getenv("SBLAS_SPLIT_CARVEOUT"); return e ? atoi(e) : -1; }();
          if (carve >= 0) cudaFuncSetAttribute((const void*)kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        }
        if ((rc = set_smem_once<T>((const void*)kern, smem)) == 0) {
          kern<<<tt->n_tiles, kSpmvBlock, smem, st>>>(tt->d_tile_row, tile, rows, nnz, rp, ci, v, x,
                                                      sink, carry_out, head_val, head_row, cap);
          ++g_launches;
          cudaError_t e = cudaGetLastError();
          if (e != cudaSuccess) { set_error("spmv split launch: %s", cudaGetErrorString(e)); rc = 3; }
        }
      }
      if (rc == 0) {
        spmv_split_fixup_kernel<T><<<(int)ceil_div(tt->n_tiles, 256), 256, 0, st>>>(
            rp, tile, tt->n_tiles, carry_out, head_val, head_row, sink);
        ++g_launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("spmv split fixup launch: %s", cudaGetErrorString(e)); rc = 3; }
      }
    }
  }
  if (rc == 0 && variant == SBLAS_SPMV_VECTOR) {
    int lanes = (o && o->lanes > 0) ? o->lanes : pow2_lanes(rows ? (double)nnz / rows : 0.0);
    if (auto_gather && !(o && o->lanes > 0)) lanes = std::min(32, lanes * 2);  // ~2 nnz per lane: more gathers in flight
    rc = (flags & SBLAS_FLAG_FMA) ? launch_vector<T, kSpmvBlock, true>(st, lanes, rows, rp, ci, v, x, sink)
                                  : launch_vector<T, kSpmvBlock, false>(st, lanes, rows, rp, ci, v, x, sink);
  }
  if (rc == 0 && variant != SBLAS_SPMV_TILE && variant != SBLAS_SPMV_SPLIT &&
      variant != SBLAS_SPMV_VECTOR && variant != SBLAS_SPMV_PIPE) {
    set_error("spmv: unknown variant %d", variant);
    rc = 2;
  }
  return rc;
}

}  // namespace sblas

using namespace sblas;

extern "C" {

int sblas_spmv_csr_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                       const int32_t* ci, const double* v, const double* x, double* y, double alpha,
                       double beta, sblasPlan_t plan, const sblasOpts* opts) {
  return spmv_impl<double>((cudaStream_t)stream, rows, cols, nnz, rp, ci, v, x, y, alpha, beta, plan, opts);
}
int sblas_spmv_csr_f32(void* stream, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                       const int32_t* ci, const float* v, const float* x, float* y, float alpha,
                       float beta, sblasPlan_t plan, const sblasOpts* opts) {
  return spmv_impl<float>((cudaStream_t)stream, rows, cols, nnz, rp, ci, v, x, y, alpha, beta, plan, opts);
}

int sblas_plan_prepare_i32(void* stream, sblasPlan_t p, const int32_t* d_colidx, int32_t cols, int32_t elem_size) {
  SBLAS_CHECK(p && (p->nnz == 0 || d_colidx) && cols >= 0 && (elem_size == 4 || elem_size == 8), "plan_prepare: bad arguments");
  if (p->rows == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  measure_gather_heavy(st, p, p->d_rowptr, d_colidx, p->rows, cols, (size_t)elem_size);
  bool ok = false;
  return prepare_pipe_table(st, p, &ok);
}

int sblas_plan_col_jump(sblasPlan_t p, double* out) {
  SBLAS_CHECK(p && out, "plan_col_jump: null");
  *out = p->col_jump;
  return 0;
}

int sblas_plan_stats(sblasPlan_t p, sblasCsrStats* out) {
  SBLAS_CHECK(p && out, "plan_stats: null");
  memset(out, 0, sizeof *out);
  out->nnz = p->nnz;
  out->rows = p->rows;
  out->max_row_nnz = p->max_row_nnz;
  out->empty_rows = p->empty_rows;
  out->spmv_tile = kSpmvDefaultTile;
  out->spmv_variant = resolve_variant<double>(p, nullptr, kSpmvDefaultTile);
  for (const TileTable& t : p->tables) {
    if (t.row_weight == 1 && t.tile == kSpmvDefaultTile) {
      out->spmv_tiles = t.n_tiles;
      out->spmv_smem_bytes = ((t.max_tile_nnz + 9) & ~3) * 8 + (t.max_tile_rows + 2) * 4;
    }
    if (t.row_weight > 1) {
      out->spmm_tile = t.tile;
      out->spmm_tiles = t.n_tiles;
    }
  }
  return 0;
}

}  // extern "C"

// ---- allocation-free SpMV (no counterpart in the reference) ------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it):
// the SPLIT kernels above (merge-path tiles that cut long rows + carry fix-up) with their scratch — the split
// table and the per-tile carry / head arrays — in a file-scope device array of fixed size instead of plan-owned
// cudaMalloc'ed buffers.  Every GPU that loads the library holds the array (3.25 MiB) until exit.  Every call shares
// it and it keeps the last call's contents: each word a call reads was written earlier in the same call (the
// split table by split_search_kernel, carry_out / head_row by every tile, head_val where head_row says so), so
// nothing has to start at zero — but two calls that overlap on the GPU (different streams, or different host
// threads) race on it.
namespace sblas {
constexpr int32_t kNoAllocTile = kSpmvSplitDefaultTile;                    // 1536 merge items per tile
constexpr int64_t kNoAllocMaxTiles = 131072;                               // rows + nnz <= 201 326 592
constexpr size_t kNoAllocScratchBytes = (3u << 20) + (256u << 10);          // 3.25 MiB: 4 + 8 + 8 + 4 B per tile, + slack
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];
static_assert((kNoAllocMaxTiles + 1) * 4 + 256 + kNoAllocMaxTiles * (8 + 8 + 4) + 256 <= (int64_t)kNoAllocScratchBytes,
              "scratch layout does not fit");

// the array's DEVICE address on the current device (the name itself is only a host-side stand-in): looked up
// once per device, with that device current
static int noalloc_scratch_ptr(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptr[64] = {nullptr};
  int dev = 0;
  SBLAS_CUDA(cudaGetDevice(&dev));
  SBLAS_CHECK(dev >= 0 && dev < 64, "spmv noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    unsigned char* p = nullptr;
    SBLAS_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    SBLAS_CHECK(p != nullptr, "spmv noalloc: cudaGetSymbolAddress returned a null address");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}
}  // namespace sblas

extern "C" int sblas_spmv_csr_noalloc_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                                          const int32_t* ci, const double* v, const double* x, double* y,
                                          double alpha, double beta) {
  using T = double;
  cudaStream_t st = (cudaStream_t)stream;
  SBLAS_CHECK(rows >= 0 && cols >= 0 && nnz >= 0 && nnz < (1ll << 31), "spmv noalloc: bad sizes");
  if (rows == 0) return 0;
  SBLAS_CHECK(rp && y && (nnz == 0 || (ci && v && x)), "spmv noalloc: null pointer");
  const int32_t tile = kNoAllocTile;
  const int64_t n_tiles64 = ceil_div(std::max<int64_t>((int64_t)rows + nnz, 1), tile);
  SBLAS_CHECK(n_tiles64 <= kNoAllocMaxTiles,
              "spmv noalloc: rows + nnz = %lld exceeds the fixed scratch (at most %lld; use sblas_spmv_csr_f64 with a plan)",
              (long long)((int64_t)rows + nnz), (long long)(kNoAllocMaxTiles * tile));
  const int32_t n_tiles = (int32_t)n_tiles64;
  unsigned char* base = nullptr;
  if (int rc = noalloc_scratch_ptr(&base)) return rc;
  int32_t* split_row = reinterpret_cast<int32_t*>(base);
  size_t off = (((size_t)kNoAllocMaxTiles + 1) * 4 + 255) & ~(size_t)255;
  T* carry_out = reinterpret_cast<T*>(base + off);
  T* head_val = carry_out + kNoAllocMaxTiles;
  int32_t* head_row = reinterpret_cast<int32_t*>(head_val + kNoAllocMaxTiles);
  RowSink<T> sink;
  sink.y = y;
  sink.out = make_outspec<T>(nullptr);
  sink.carry = nullptr;
  sink.rows = rows;
  sink.flags = 0;
  sink.alpha = alpha;
  sink.beta = beta;
  sink.part = nullptr;
  sink.stage_src = y;
  sink.phase = 0;
  split_search_kernel<<<(int)ceil_div(n_tiles + 1, 256), 256, 0, st>>>(rp, rows, nnz, tile, n_tiles, split_row);
  SBLAS_LAUNCH_CHECK();
  const int32_t cap = (tile + 9) & ~3;
  const int smem = cap * (int)sizeof(T) + (tile + 3) * (int)sizeof(int32_t) + ((tile + 31) / 32 + 2) * 4;  // < 48 KB: no attribute call
  if (aligned16(ci) && aligned16(v))
    spmv_split_kernel<T, kSpmvBlock, true, true><<<n_tiles, kSpmvBlock, smem, st>>>(split_row, tile, rows, nnz, rp, ci, v, x, sink,
                                                                                   carry_out, head_val, head_row, cap);
  else
    spmv_split_kernel<T, kSpmvBlock, false, true><<<n_tiles, kSpmvBlock, smem, st>>>(split_row, tile, rows, nnz, rp, ci, v, x, sink,
                                                                                    carry_out, head_val, head_row, cap);
  SBLAS_LAUNCH_CHECK();
  spmv_split_fixup_kernel<T><<<(int)ceil_div(n_tiles, 256), 256, 0, st>>>(rp, tile, n_tiles, carry_out, head_val, head_row, sink);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
