// spmv.cu — CSR SpMV kernels for sm_100a:  y <- beta*y + alpha*A*x.
//
// Replaces cusparseSpMV (spmv.h:96-106) and the separate axpby pass (spmv.h:138,
// kernel.h:28-37) of the reference; the CPU statement of the same arithmetic is
// sblas_spmv_csr_cpu (spmv.h:16-32).
//
// Three variants (sblasOpts.variant, AUTO picks from the plan's row statistics):
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
#include "plan.cuh"

namespace sblas {

constexpr int kSpmvBlock = 256;
constexpr int kSpmvDefaultTile = 3072;
constexpr int kSpmvMaxSmem = 96 * 1024;

template <typename T>
struct RowSink {
  T* y;
  OutSpec<T> out;
  T* carry;       // head partial sum at [0], tail partial sum at [4] (= kpad of k = 1, sblas_b200.h)
  int32_t rows;   // local row count
  int32_t flags;  // HEAD_PARTIAL | TAIL_PARTIAL
  T alpha, beta;

  __device__ __forceinline__ void put(int32_t r, T sum) const {
    if ((flags & SBLAS_FLAG_HEAD_PARTIAL) && r == 0) { carry[0] = sum; return; }
    if ((flags & SBLAS_FLAG_TAIL_PARTIAL) && r == rows - 1) { carry[4] = sum; return; }
    T res = axpby_rn(alpha, sum, beta, y[r]);
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
template <typename T, int BLOCK, bool VEC>
__device__ __forceinline__ void stage_products(const int32_t* __restrict__ colidx,
                                               const T* __restrict__ val,
                                               const T* __restrict__ x, int32_t n0, int32_t n1,
                                               T* s_prod) {
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
        p[0] = mul_rn(v[0], __ldg(x + c.x));
        p[1] = mul_rn(v[1], __ldg(x + c.y));
        p[2] = mul_rn(v[2], __ldg(x + c.z));
        p[3] = mul_rn(v[3], __ldg(x + c.w));
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = base + e;
          p[e] = (j >= n0 && j < n1) ? mul_rn(ldg_stream(val + j), __ldg(x + ldg_stream(colidx + j)))
                                     : T(0);
        }
      }
      store4_smem(s_prod + (base - n0a), p);
    }
  } else {
#pragma unroll 4
    for (int j = n0 + tid; j < n1; j += BLOCK)
      s_prod[j - n0a] = mul_rn(ldg_stream(val + j), __ldg(x + ldg_stream(colidx + j)));
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

template <typename T, int BLOCK, bool VEC>
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
  // local row pointers clamped to the tile's nnz window; entry nfin+1 closes the cut row rb
  for (int i = tid; i <= nfin; i += BLOCK) {
    int32_t v = rowptr[ra + i];
    s_rp[i] = v < ja ? ja : (v > jb ? jb : v);
  }
  if (tid == 0) s_rp[nfin + 1] = jb;
  __syncthreads();
  const bool head_open = nfin > 0 && rowptr[ra] < ja;
  if (tid == 0) head_row[b] = head_open ? ra : -1;
  stage_products<T, BLOCK, VEC>(colidx, val, x, ja, jb, s_prod);
  __syncthreads();
  const int nloc = rb < rows ? nfin + 1 : nfin;  // + the cut row, if there is one
  if (rb >= rows && tid == 0) carry_out[b] = T(0);
  if (nloc <= 0) return;
  const SplitSink<T> ss{sink, ra, rb < rows ? nfin : -1, head_open, carry_out + b, head_val + b};
  reduce_rows_dispatch<T, BLOCK>(lanes_for_avg((jb - ja) / nloc), s_prod, s_rp, nloc, ja & ~3, ss);
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

template <typename T>
static int resolve_variant(const sblasPlan_st* plan, const sblasOpts* o, int32_t tile) {
  if (o && o->variant != SBLAS_VARIANT_AUTO) return o->variant;
  if (!plan) return SBLAS_SPMV_VECTOR;
  return plan->max_row_nnz <= tile / 2 ? SBLAS_SPMV_TILE : SBLAS_SPMV_SPLIT;
}

template <typename T>
static int set_smem_once(const void* fn, int bytes) {
  if (bytes > 48 * 1024) {
    SBLAS_CHECK(bytes <= 227 * 1024, "spmv: tile needs %d bytes of shared memory", bytes);
    SBLAS_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  }
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
  const int32_t tile = (o && o->tile > 0) ? o->tile : kSpmvDefaultTile;
  sblasPlan_st* eph = nullptr;
  int variant = resolve_variant<T>(plan, o, tile);
  if (!plan && variant != SBLAS_SPMV_VECTOR) {
    if (sblas_plan_create_i32(st, rows, nnz, rp, &eph)) return 4;
    plan = eph;
  }
  if (plan) SBLAS_CHECK(plan->rows == rows && plan->nnz == nnz, "spmv: plan does not match matrix");
  const bool vec = aligned16(ci) && aligned16(v);
  int rc = 0;
  if (variant == SBLAS_SPMV_TILE) {
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
    // table (row_weight = -1) marks the merge-path split table; built here, cached in the plan
    const TileTable* tt = nullptr;
    for (const TileTable& t : plan->tables)
      if (t.tile == tile && t.row_weight == -1) tt = &t;
    if (!tt) {
      TileTable t;
      t.tile = tile;
      t.row_weight = -1;
      t.n_tiles = (int32_t)ceil_div(std::max<int64_t>((int64_t)rows + nnz, 1), tile);
      SBLAS_CUDA(cudaMalloc(&t.d_tile_row, sizeof(int32_t) * (size_t)(t.n_tiles + 1)));
      split_search_kernel<<<(int)ceil_div(t.n_tiles + 1, 256), 256, 0, st>>>(rp, rows, nnz, tile,
                                                                           t.n_tiles, t.d_tile_row);
      SBLAS_LAUNCH_CHECK();
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
      const int32_t cap = (tile + 9) & ~3;
      const int smem = cap * (int)sizeof(T) + (tile + 3) * (int)sizeof(int32_t);
      auto kern = vec ? spmv_split_kernel<T, kSpmvBlock, true> : spmv_split_kernel<T, kSpmvBlock, false>;
      if ((rc = set_smem_once<T>((const void*)kern, smem)) == 0) {
        kern<<<tt->n_tiles, kSpmvBlock, smem, st>>>(tt->d_tile_row, tile, rows, nnz, rp, ci, v, x,
                                                    sink, carry_out, head_val, head_row, cap);
        spmv_split_fixup_kernel<T><<<(int)ceil_div(tt->n_tiles, 256), 256, 0, st>>>(
            rp, tile, tt->n_tiles, carry_out, head_val, head_row, sink);
        g_launches += 2;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("spmv split launch: %s", cudaGetErrorString(e)); rc = 3; }
      }
    }
  }
  if (rc == 0 && variant == SBLAS_SPMV_VECTOR) {
    int lanes = (o && o->lanes > 0) ? o->lanes : pow2_lanes(rows ? (double)nnz / rows : 0.0);
    rc = (flags & SBLAS_FLAG_FMA) ? launch_vector<T, kSpmvBlock, true>(st, lanes, rows, rp, ci, v, x, sink)
                                  : launch_vector<T, kSpmvBlock, false>(st, lanes, rows, rp, ci, v, x, sink);
  }
  if (rc == 0 && variant != SBLAS_SPMV_TILE && variant != SBLAS_SPMV_SPLIT &&
      variant != SBLAS_SPMV_VECTOR) {
    set_error("spmv: unknown variant %d", variant);
    rc = 2;
  }
  if (eph) {
    cudaStreamSynchronize(st);
    sblas_plan_destroy(eph);
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
