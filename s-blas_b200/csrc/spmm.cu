// spmm.cu — CSR SpMM kernels for sm_100a:  C <- beta*C + alpha*A*B, k dense columns.
//
// Replaces cusparseSpMM (spmm.h:135-149 method 1, spmm.h:240-251 method 2) and the separate
// axpby pass (spmm.h:283, kernel.h:28-37); CPU statement: sblas_spmm_csr_cpu (spmm.h:30-70).
//
//  ROWMAJOR  The internal device layout of B and C is row-major (DenseMatrix::sync2gpu
//            transposes; see include/matrix.h), so the B row selected by one nnz is one
//            contiguous k*sizeof(T) segment.  A group of G lanes owns one A row: it loads 32
//            (col,val) pairs coalesced, broadcasts them with warp shuffles, and every lane
//            gathers its 16-byte slice of the B row (LDG.128, fully coalesced across the
//            group) and keeps NB vector accumulators in registers, so col_idx/val are read
//            ONCE for all k columns.  Each (row, column) sum is accumulated by one lane in
//            CSR order with separately rounded mul/add, so the result is bit-identical to
//            the reference CPU loop (SBLAS_FLAG_FMA switches to fused multiply-add).
//            The beta/alpha epilogue and the stores into every GPU's replica of C are fused.
//            Bound: HBM (gather model: one k*sizeof(T) B-row segment per nnz when B does not
//            fit in L2; L1/L2 reuse for banded inputs) — see DESIGN.md.
//  COLMAJOR  Native col-major B/C (the reference's host layout, spmm.h:121-132): a tile of A
//            is staged once in shared memory and reused for all k columns ("k SpMVs sharing
//            A").  Used by the managed-memory variant (no copies, matrix_um.h) and for parity.
#include "plan.cuh"

namespace sblas {

constexpr int kSpmmBlock = 256;
constexpr int kSpmmDefaultTile = 2048;
constexpr int kSpmmRowWeight = 2;

template <typename T>
struct MatSink {
  T* C;
  int64_t ldc;
  OutSpec<T> out;
  T* carry;       // [2][kpad] head / tail partial sums (raw, before alpha/beta)
  int32_t kpad;
  int32_t rows;
  int32_t k;
  int32_t flags;
  T alpha, beta;

  __device__ __forceinline__ int partial_slot(int32_t r) const {
    if ((flags & SBLAS_FLAG_HEAD_PARTIAL) && r == 0) return 0;
    if ((flags & SBLAS_FLAG_TAIL_PARTIAL) && r == rows - 1) return 1;
    return -1;
  }
};

// ---- small vector helpers -------------------------------------------------------------------
template <typename T, int VEC> struct VecT;
template <> struct VecT<double, 2> { using type = double2; };
template <> struct VecT<float, 4> { using type = float4; };
template <> struct VecT<double, 1> { using type = double; };
template <> struct VecT<float, 1> { using type = float; };

__device__ __forceinline__ void vzero(double2& a) { a.x = a.y = 0.0; }
__device__ __forceinline__ void vzero(float4& a) { a.x = a.y = a.z = a.w = 0.f; }
__device__ __forceinline__ void vzero(double& a) { a = 0.0; }
__device__ __forceinline__ void vzero(float& a) { a = 0.f; }

template <bool EXACT, typename T>
__device__ __forceinline__ T mac(T v, T b, T acc) {
  return EXACT ? add_rn(acc, mul_rn(v, b)) : fma_rn(v, b, acc);
}
template <bool EXACT> __device__ __forceinline__ void vmac(double v, const double2& b, double2& a) {
  a.x = mac<EXACT>(v, b.x, a.x); a.y = mac<EXACT>(v, b.y, a.y);
}
template <bool EXACT> __device__ __forceinline__ void vmac(float v, const float4& b, float4& a) {
  a.x = mac<EXACT>(v, b.x, a.x); a.y = mac<EXACT>(v, b.y, a.y);
  a.z = mac<EXACT>(v, b.z, a.z); a.w = mac<EXACT>(v, b.w, a.w);
}
template <bool EXACT> __device__ __forceinline__ void vmac(double v, const double& b, double& a) { a = mac<EXACT>(v, b, a); }
template <bool EXACT> __device__ __forceinline__ void vmac(float v, const float& b, float& a) { a = mac<EXACT>(v, b, a); }

__device__ __forceinline__ double2 vaxpby(double al, const double2& s, double be, const double2& c) {
  return make_double2(axpby_rn(al, s.x, be, c.x), axpby_rn(al, s.y, be, c.y));
}
__device__ __forceinline__ float4 vaxpby(float al, const float4& s, float be, const float4& c) {
  return make_float4(axpby_rn(al, s.x, be, c.x), axpby_rn(al, s.y, be, c.y),
                     axpby_rn(al, s.z, be, c.z), axpby_rn(al, s.w, be, c.w));
}
__device__ __forceinline__ double vaxpby(double al, const double& s, double be, const double& c) { return axpby_rn(al, s, be, c); }
__device__ __forceinline__ float vaxpby(float al, const float& s, float be, const float& c) { return axpby_rn(al, s, be, c); }

__device__ __forceinline__ double ldg_v(const double* p) { return __ldg(p); }
__device__ __forceinline__ float ldg_v(const float* p) { return __ldg(p); }
__device__ __forceinline__ double2 vload_b(const double2* p, bool nol1) { return nol1 ? ldg_stream_d2(p) : __ldg(p); }
__device__ __forceinline__ float4 vload_b(const float4* p, bool nol1) { return nol1 ? ldg_stream_f4(p) : __ldg(p); }
__device__ __forceinline__ double vload_b(const double* p, bool nol1) { return nol1 ? ldg_stream(p) : __ldg(p); }
__device__ __forceinline__ float vload_b(const float* p, bool nol1) { return nol1 ? ldg_stream(p) : __ldg(p); }

// ---- ROWMAJOR ---------------------------------------------------------------------------------
// G lanes per A row, VEC elements per lane per column block, NB column blocks per lane:
// covers k <= G*VEC*NB columns (larger k: the host loops over column panels).
template <typename T, int VEC, int G, int NB, bool EXACT, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
spmm_rm_kernel(const int32_t* __restrict__ tile_row, const int32_t* __restrict__ rowptr,
               const int32_t* __restrict__ colidx, const T* __restrict__ val,
               const T* __restrict__ B, int64_t ldb, MatSink<T> sink) {
  using V = typename VecT<T, VEC>::type;
  constexpr int GPB = BLOCK / G;  // row groups per CTA
  constexpr int U = NB >= 8 ? 1 : NB >= 4 ? 2 : 4;  // nnz per step (MLP = U*NB gathers per lane)
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const int gi = threadIdx.x / G;
  const int32_t r0 = tile_row[blockIdx.x], r1 = tile_row[blockIdx.x + 1];
  const int k = sink.k;

  for (int32_t rb = r0; rb < r1; rb += GPB) {  // block-uniform
    const int32_t r = rb + gi;
    const bool valid = r < r1;
    const int32_t s = valid ? rowptr[r] : 0;
    const int32_t len = valid ? rowptr[r + 1] - s : 0;
    const int32_t maxlen = (G == 32) ? len : warp_max(len);
    V acc[NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) vzero(acc[nb]);

    // column guards are loop-invariant per lane
    bool colok[NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) colok[nb] = (nb * G + gl) * VEC < k;
    const T* bbase = B + gl * VEC;

    for (int32_t j0 = 0; j0 < maxlen; j0 += G) {
      int32_t c = 0;
      T v = T(0);
      if (j0 + gl < len) {
        c = ldg_stream(colidx + s + j0 + gl);
        v = ldg_stream(val + s + j0 + gl);
      }
      const int cnt = min(G, (int)(len - j0));            // this group's nnz in the chunk
      const int wcnt = min(G, (int)(maxlen - j0));        // warp-uniform
      // U nnz per step: all their B-row loads are issued before the first multiply-add, so
      // every lane keeps U*NB independent 16-byte gathers in flight.
      for (int t0 = 0; t0 < wcnt; t0 += U) {
        V bv[U][NB];
        T vv[U];
        bool act[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int t = t0 + u;
          const int32_t cc = __shfl_sync(0xffffffffu, c, t & (G - 1), G);
          vv[u] = __shfl_sync(0xffffffffu, v, t & (G - 1), G);
          act[u] = t < cnt;
          const T* brow = bbase + (int64_t)cc * ldb;
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            vzero(bv[u][nb]);
            if (act[u] && colok[nb]) bv[u][nb] = ldg_v(reinterpret_cast<const V*>(brow + nb * G * VEC));
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (act[u]) {
#pragma unroll
            for (int nb = 0; nb < NB; ++nb)
              if (colok[nb]) vmac<EXACT>(vv[u], bv[u][nb], acc[nb]);
          }
        }
      }
    }

    if (valid) {
      const int slot = sink.partial_slot(r);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const int c0 = (nb * G + gl) * VEC;
        if (c0 < k) {
          if (slot >= 0) {
            *reinterpret_cast<V*>(sink.carry + (int64_t)slot * sink.kpad + c0) = acc[nb];
          } else {
            const int64_t off = (int64_t)r * sink.ldc + c0;
            const V cv = *reinterpret_cast<const V*>(sink.C + off);
            const V res = vaxpby(sink.alpha, acc[nb], sink.beta, cv);
            *reinterpret_cast<V*>(sink.C + off) = res;
            for (int d = 0; d < sink.out.n; ++d) *reinterpret_cast<V*>(sink.out.ptr[d] + off) = res;
          }
        }
      }
    }
  }
}

// ---- ROWBLOCK (row-major B/C, R x 1 BCSR mirror of A, spmm_rb.cu) --------------------------------
// One warp per block row (R adjacent rows); lane l owns columns (nb*32 + l)*VEC .. +VEC of the
// panel for nb < NB.  Per block: one uniform (broadcast) load of {col, mask} and of the R values,
// ONE gather of the B row slice, and up to R*NB vector multiply-adds from registers: the B slice
// is reused R-fold.  U blocks are in flight per step.
template <typename T, int VEC, int NB, int R, bool EXACT, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
spmm_rb_kernel(int64_t n_blockrows, const int32_t* __restrict__ blk_ptr, const int2* __restrict__ colmask,
               const T* __restrict__ bval, const T* __restrict__ B, int64_t ldb, MatSink<T> sink) {
  using V = typename VecT<T, VEC>::type;
  constexpr int U = NB >= 4 ? 1 : NB == 2 ? 2 : 4;
  constexpr int WPB = BLOCK / 32;
  const int lane = threadIdx.x & 31;
  const int k = sink.k;
  bool colok[NB];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) colok[nb] = (nb * 32 + lane) * VEC < k;
  const T* bbase = B + lane * VEC;
  for (int64_t br = (int64_t)blockIdx.x * WPB + (threadIdx.x >> 5); br < n_blockrows; br += (int64_t)gridDim.x * WPB) {
    const int32_t b0 = blk_ptr[br], b1 = blk_ptr[br + 1];
    V acc[R][NB];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) vzero(acc[r][nb]);
    for (int32_t b = b0; b < b1; b += U) {
      int2 cm[U];
      T vv[U][R];
      V bv[U][NB];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        cm[u] = make_int2(0, 0);
        if (b + u < b1) {
          cm[u] = __ldg(colmask + b + u);
#pragma unroll
          for (int r = 0; r < R; ++r) vv[u][r] = __ldg(bval + (int64_t)(b + u) * R + r);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const T* brow = bbase + (int64_t)cm[u].x * ldb;
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          vzero(bv[u][nb]);
          if (cm[u].y != 0 && colok[nb]) bv[u][nb] = ldg_v(reinterpret_cast<const V*>(brow + nb * 32 * VEC));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (cm[u].y & (1 << r)) {
#pragma unroll
            for (int nb = 0; nb < NB; ++nb)
              if (colok[nb]) vmac<EXACT>(vv[u][r], bv[u][nb], acc[r][nb]);
          }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int64_t row = br * R + r;
      if (row >= sink.rows) break;
      const int slot = sink.partial_slot((int32_t)row);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const int c0 = (nb * 32 + lane) * VEC;
        if (c0 < k) {
          if (slot >= 0) {
            *reinterpret_cast<V*>(sink.carry + (int64_t)slot * sink.kpad + c0) = acc[r][nb];
          } else {
            const int64_t off = row * sink.ldc + c0;
            const V cv = *reinterpret_cast<const V*>(sink.C + off);
            const V res = vaxpby(sink.alpha, acc[r][nb], sink.beta, cv);
            *reinterpret_cast<V*>(sink.C + off) = res;
            for (int d = 0; d < sink.out.n; ++d) *reinterpret_cast<V*>(sink.out.ptr[d] + off) = res;
          }
        }
      }
    }
  }
}

template <typename T, int VEC, bool EXACT>
static int launch_rb(cudaStream_t st, const sblasPlan_st* plan, int NB, const T* B, int64_t ldb,
                     const MatSink<T>& sink) {
  const RowBlocks& rb = plan->rb;
  constexpr int BLOCK = 256;
  const int grid = (int)std::min<int64_t>(ceil_div(rb.n_blockrows, BLOCK / 32), (int64_t)kNumSM * 64);
#define SBLAS_RB(NN)                                                                              \
  spmm_rb_kernel<T, VEC, NN, 4, EXACT, BLOCK><<<grid, BLOCK, 0, st>>>(                             \
      rb.n_blockrows, rb.d_ptr, (const int2*)rb.d_colmask, (const T*)rb.d_val, B, ldb, sink)
  if (NB <= 1) SBLAS_RB(1);
  else if (NB == 2) SBLAS_RB(2);
  else SBLAS_RB(4);
#undef SBLAS_RB
  SBLAS_LAUNCH_CHECK();
  return 0;
}

// ---- COLMAJOR ---------------------------------------------------------------------------------
// Tile of A staged in shared memory once, then one pass per dense column: G lanes per row read
// consecutive (col,val) pairs from shared memory (conflict-free), gather B[n*ldb + col] through
// L1 and reduce with warp shuffles.  G == 1 accumulates sequentially (bit-identical to
// spmm.h:56-68).  Warps take columns round-robin, so the C stores of a warp are contiguous runs
// of 32/G rows in one column.
template <typename T, int G, int BLOCK, bool EXACT>
__device__ __forceinline__ void cm_columns(const int32_t* s_col, const T* s_val,
                                           const int32_t* s_rp, int32_t nrows, int32_t n0,
                                           int32_t r0, const T* __restrict__ B, int64_t ldb,
                                           const MatSink<T>& sink) {
  constexpr int NW = BLOCK / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gl = lane & (G - 1), gw = lane / G;
  constexpr int RPW = 32 / G;  // rows per warp pass
  for (int n = warp; n < sink.k; n += NW) {
    const T* bcol = B + (int64_t)n * ldb;
    for (int i0 = 0; i0 < nrows; i0 += RPW) {  // warp-uniform
      const int i = i0 + gw;
      const bool valid = i < nrows;
      const int s = valid ? s_rp[i] - n0 : 0, e = valid ? s_rp[i + 1] - n0 : 0;
      T sum = 0;
      for (int j = s + gl; j < e; j += G) sum = mac<EXACT>(s_val[j], __ldg(bcol + s_col[j]), sum);
#pragma unroll
      for (int m = G / 2; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
      if (valid && gl == 0) {
        const int32_t r = r0 + i;
        const int slot = sink.partial_slot(r);
        if (slot >= 0) {
          sink.carry[(int64_t)slot * sink.kpad + n] = sum;
        } else {
          const int64_t off = (int64_t)n * sink.ldc + r;
          const T res = axpby_rn(sink.alpha, sum, sink.beta, sink.C[off]);
          sink.C[off] = res;
          for (int d = 0; d < sink.out.n; ++d) sink.out.ptr[d][off] = res;
        }
      }
    }
  }
}

template <typename T, int BLOCK, bool EXACT>
__global__ void __launch_bounds__(BLOCK)
spmm_cm_kernel(const int32_t* __restrict__ tile_row, const int32_t* __restrict__ rowptr,
               const int32_t* __restrict__ colidx, const T* __restrict__ val,
               const T* __restrict__ B, int64_t ldb, MatSink<T> sink, int32_t cap_nnz,
               int32_t force_lanes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* s_val = reinterpret_cast<T*>(smem_raw);
  int32_t* s_col = reinterpret_cast<int32_t*>(s_val + cap_nnz);
  int32_t* s_rp = s_col + cap_nnz;
  const int tid = threadIdx.x;
  const int32_t r0 = tile_row[blockIdx.x], r1 = tile_row[blockIdx.x + 1];
  const int32_t nrows = r1 - r0;
  if (nrows <= 0) return;
  for (int i = tid; i <= nrows; i += BLOCK) s_rp[i] = rowptr[r0 + i];
  __syncthreads();
  const int32_t n0 = s_rp[0], n1 = s_rp[nrows];
  for (int j = n0 + tid; j < n1; j += BLOCK) {
    s_col[j - n0] = ldg_stream(colidx + j);
    s_val[j - n0] = ldg_stream(val + j);
  }
  __syncthreads();
  int lanes = force_lanes;
  if (lanes == 0) {
    const int avg = (n1 - n0) / nrows;
    lanes = avg <= 6 ? 1 : avg <= 12 ? 2 : avg <= 24 ? 4 : avg <= 48 ? 8 : avg <= 96 ? 16 : 32;
  }
  switch (lanes) {
    case 1: cm_columns<T, 1, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
    case 2: cm_columns<T, 2, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
    case 4: cm_columns<T, 4, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
    case 8: cm_columns<T, 8, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
    case 16: cm_columns<T, 16, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
    default: cm_columns<T, 32, BLOCK, EXACT>(s_col, s_val, s_rp, nrows, n0, r0, B, ldb, sink); break;
  }
}

// ---- host side ---------------------------------------------------------------------------------
template <typename T, int VEC, bool EXACT>
static int launch_rm(cudaStream_t st, const TileTable* tt, int G, int NB, const int32_t* rp,
                     const int32_t* ci, const T* v, const T* B, int64_t ldb,
                     const MatSink<T>& sink) {
#define SBLAS_RM(GG, NN)                                                                         \
  spmm_rm_kernel<T, VEC, GG, NN, EXACT, kSpmmBlock>                                              \
      <<<tt->n_tiles, kSpmmBlock, 0, st>>>(tt->d_tile_row, rp, ci, v, B, ldb, sink)
  if (G == 1) SBLAS_RM(1, 1);
  else if (G == 2) SBLAS_RM(2, 1);
  else if (G == 4) SBLAS_RM(4, 1);
  else if (G == 8) SBLAS_RM(8, 1);
  else if (G == 16) SBLAS_RM(16, 1);
  else if (NB == 1) SBLAS_RM(32, 1);
  else if (NB == 2) SBLAS_RM(32, 2);
  else if (NB <= 4) SBLAS_RM(32, 4);
  else SBLAS_RM(32, 8);
#undef SBLAS_RM
  SBLAS_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int spmm_impl(cudaStream_t st, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                     const int32_t* ci, const T* v, int32_t k, int32_t layout, const T* B,
                     int64_t ldb, T* C, int64_t ldc, T alpha, T beta, sblasPlan_st* plan,
                     const sblasOpts* o) {
  SBLAS_CHECK(rows >= 0 && cols >= 0 && nnz >= 0 && nnz < (1ll << 31) && k >= 0, "spmm: bad sizes");
  if (rows == 0 || k == 0) return 0;
  SBLAS_CHECK(rp && C && (nnz == 0 || (ci && v && B)), "spmm: null pointer");
  SBLAS_CHECK(layout == SBLAS_ROW_MAJOR || layout == SBLAS_COL_MAJOR, "spmm: bad layout %d", layout);
  const int flags = o ? o->flags : 0;
  SBLAS_CHECK(!(flags & (SBLAS_FLAG_HEAD_PARTIAL | SBLAS_FLAG_TAIL_PARTIAL)) || (o && o->carry),
              "spmm: HEAD/TAIL_PARTIAL needs opts->carry");
  if (layout == SBLAS_ROW_MAJOR)
    SBLAS_CHECK(ldb >= k && ldc >= k, "spmm: row-major ld must be >= k");
  else
    SBLAS_CHECK(ldb >= cols && ldc >= rows, "spmm: col-major ld must be >= rows of the operand");

  sblasPlan_st* eph = nullptr;
  if (!plan) {
    if (sblas_plan_create_i32(st, rows, nnz, rp, &eph)) return 4;
    plan = eph;
  }
  SBLAS_CHECK(plan->rows == rows && plan->nnz == nnz, "spmm: plan does not match matrix");
  const int32_t tile = (o && o->tile > 0) ? o->tile : kSpmmDefaultTile;
  const TileTable* tt = nullptr;
  int rc = plan_get_table(plan, st, kSpmmRowWeight, tile, &tt);

  MatSink<T> sink;
  sink.C = C;
  sink.ldc = ldc;
  sink.out = make_outspec<T>(o);
  sink.carry = o ? (T*)o->carry : nullptr;
  sink.kpad = (k + 3) & ~3;
  sink.rows = rows;
  sink.k = k;
  sink.flags = flags;
  sink.alpha = alpha;
  sink.beta = beta;
  const bool exact = !(flags & SBLAS_FLAG_FMA);

  if (rc == 0 && layout == SBLAS_ROW_MAJOR) {
    constexpr int VF = Vec16<T>::N;
    bool vec = aligned16(B) && aligned16(C) && (ldb % VF == 0) && (ldc % VF == 0) &&
               ldb >= ((k + VF - 1) / VF) * VF && ldc >= ((k + VF - 1) / VF) * VF &&
               (sink.carry == nullptr || aligned16(sink.carry));
    for (int d = 0; d < sink.out.n; ++d) vec = vec && aligned16(sink.out.ptr[d]);
    const int VEC = vec ? VF : 1;
    // row-block mirror (R x 1 BCSR, built by sblas_plan_build_rowblocks_*): one B load feeds up to
    // R adjacent rows.  Taken when the plan holds it for this value type, the operands are
    // vectorisable and a row slice spans a whole warp (k >= 32*VEC); AUTO or SBLAS_SPMM_ROWBLOCK.
    const int want = o ? o->variant : SBLAS_VARIANT_AUTO;
    const bool rb_ok = vec && plan->rb.n_blocks > 0 && plan->rb.is_f64 == (sizeof(T) == 8) && k >= 32 * VF;
    if (want == SBLAS_SPMM_ROWBLOCK && !rb_ok) {
      set_error("spmm: SBLAS_SPMM_ROWBLOCK needs a plan with row blocks of this value type, 16-byte aligned operands and k >= %d", 32 * VF);
      rc = 2;
    }
    if (rc == 0 && rb_ok && (want == SBLAS_VARIANT_AUTO || want == SBLAS_SPMM_ROWBLOCK)) {
      const int rbpanel = 32 * VF * 4;
      for (int32_t c_off = 0; c_off < k && rc == 0; c_off += rbpanel) {
        const int kp = std::min<int32_t>(rbpanel, k - c_off);
        const int NB = (kp + 32 * VF - 1) / (32 * VF);
        MatSink<T> ps = sink;
        ps.C = C + c_off;
        ps.k = kp;
        for (int d = 0; d < ps.out.n; ++d) ps.out.ptr[d] = sink.out.ptr[d] + c_off;
        if (ps.carry) ps.carry = sink.carry + c_off;
        rc = exact ? launch_rb<T, VF, true>(st, plan, NB, B + c_off, ldb, ps)
                   : launch_rb<T, VF, false>(st, plan, NB, B + c_off, ldb, ps);
      }
      if (eph) { cudaStreamSynchronize(st); sblas_plan_destroy(eph); }
      return rc;
    }
    const int panel = 32 * VEC * 8;  // columns one launch covers at most
    for (int32_t c_off = 0; c_off < k && rc == 0; c_off += panel) {
      const int kp = std::min<int32_t>(panel, k - c_off);
      const int slots = (kp + VEC - 1) / VEC;  // lane-slots needed
      int G = 1;
      while (G < 32 && G < slots) G <<= 1;
      if (o && o->lanes > 0 && o->lanes >= G && o->lanes <= 32) G = o->lanes;
      const int NB = (slots + G - 1) / G;
      MatSink<T> ps = sink;
      ps.C = C + c_off;
      ps.k = kp;
      for (int d = 0; d < ps.out.n; ++d) ps.out.ptr[d] = sink.out.ptr[d] + c_off;
      if (ps.carry) ps.carry = sink.carry + c_off;
      if (vec)
        rc = exact ? launch_rm<T, VF, true>(st, tt, G, NB, rp, ci, v, B + c_off, ldb, ps)
                   : launch_rm<T, VF, false>(st, tt, G, NB, rp, ci, v, B + c_off, ldb, ps);
      else
        rc = exact ? launch_rm<T, 1, true>(st, tt, G, NB, rp, ci, v, B + c_off, ldb, ps)
                   : launch_rm<T, 1, false>(st, tt, G, NB, rp, ci, v, B + c_off, ldb, ps);
    }
  } else if (rc == 0) {
    const int32_t cap = tt->max_tile_nnz + 1;
    const int smem = cap * (int)(sizeof(T) + sizeof(int32_t)) + (tt->max_tile_rows + 2) * (int)sizeof(int32_t);
    if (smem > 200 * 1024) {
      set_error("spmm col-major: a row of %d nnz does not fit the shared-memory tile", plan->max_row_nnz);
      rc = 5;
    } else {
      auto kern = exact ? spmm_cm_kernel<T, kSpmmBlock, true> : spmm_cm_kernel<T, kSpmmBlock, false>;
      if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute((const void*)kern,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("spmm: smem attribute: %s", cudaGetErrorString(e)); rc = 1; }
      }
      if (rc == 0) {
        kern<<<tt->n_tiles, kSpmmBlock, smem, st>>>(tt->d_tile_row, rp, ci, v, B, ldb, sink, cap,
                                                    o ? o->lanes : 0);
        ++g_launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("spmm cm launch: %s", cudaGetErrorString(e)); rc = 3; }
      }
    }
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
int sblas_spmm_csr_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                       const int32_t* ci, const double* v, int32_t k, int32_t layout,
                       const double* B, int64_t ldb, double* C, int64_t ldc, double alpha,
                       double beta, sblasPlan_t plan, const sblasOpts* opts) {
  return spmm_impl<double>((cudaStream_t)stream, rows, cols, nnz, rp, ci, v, k, layout, B, ldb, C,
                           ldc, alpha, beta, plan, opts);
}
int sblas_spmm_csr_f32(void* stream, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rp,
                       const int32_t* ci, const float* v, int32_t k, int32_t layout, const float* B,
                       int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                       sblasPlan_t plan, const sblasOpts* opts) {
  return spmm_impl<float>((cudaStream_t)stream, rows, cols, nnz, rp, ci, v, k, layout, B, ldb, C,
                          ldc, alpha, beta, plan, opts);
}
}  // extern "C"
