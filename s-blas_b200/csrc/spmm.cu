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

__device__ __forceinline__ double vget(const double2& a, int e) { return e == 0 ? a.x : a.y; }
__device__ __forceinline__ float vget(const float4& a, int e) { return e == 0 ? a.x : e == 1 ? a.y : e == 2 ? a.z : a.w; }
__device__ __forceinline__ double vget(const double& a, int) { return a; }
__device__ __forceinline__ float vget(const float& a, int) { return a; }

// Fused epilogue of one lane: result columns [c0, c0 + VEC) of the row at element offset `off`.  A row's last
// vector may be partial (k not a multiple of VEC): only columns < k are read and written then, so whatever the
// caller keeps in C[.., k .. ld) survives (a sub-matrix view, or the padding of DenseMatrix's device mirror).
template <typename T, int VEC>
__device__ __forceinline__ void emit_row_slice(const MatSink<T>& sink, int64_t off, int c0,
                                               const typename VecT<T, VEC>::type& acc) {
  using V = typename VecT<T, VEC>::type;
  if (VEC == 1 || c0 + VEC <= sink.k) {
    const V cv = *reinterpret_cast<const V*>(sink.C + off);
    const V res = vaxpby(sink.alpha, acc, sink.beta, cv);
    *reinterpret_cast<V*>(sink.C + off) = res;
    for (int d = 0; d < sink.out.n; ++d) *reinterpret_cast<V*>(sink.out.ptr[d] + off) = res;
  } else {
#pragma unroll
    for (int e = 0; e < VEC; ++e)
      if (c0 + e < sink.k) {
        const T res = axpby_rn(sink.alpha, vget(acc, e), sink.beta, sink.C[off + e]);
        sink.C[off + e] = res;
        for (int d = 0; d < sink.out.n; ++d) sink.out.ptr[d][off + e] = res;
      }
  }
}

// ---- ROWMAJOR ---------------------------------------------------------------------------------
// G lanes per A row, VEC elements per lane per column block, NB column blocks per lane:
// covers k <= G*VEC*NB columns (larger k: the host loops over column panels).
// BULK (the stitch of method 2 / SpMV v1 for wide rows): the finished rows of a chunk of the tile are also parked
// in shared memory — consecutive rows, so with ldc == k one contiguous block of C — and pushed to every peer
// replica by TMA bulk stores (cp.async.bulk.global.shared::cta, SASS UBLKCP: one instruction per peer and chunk,
// 16-32 KB each) instead of seven 16-byte st.global per lane: the LSU of the computing warps carries no NVLink
// traffic, the copies drain while the CTA's other rows (and the SM's other CTAs) compute.  chunk_rows (a
// multiple of GPB) rows fit the dynamic shared memory; rows that go to the carry buffer are left out.
template <typename T, int VEC, int G, int NB, bool EXACT, int BLOCK, bool BULK = false>
__global__ void __launch_bounds__(BLOCK)
spmm_rm_kernel(const int32_t* __restrict__ tile_row, const int32_t* __restrict__ rowptr,
               const int32_t* __restrict__ colidx, const T* __restrict__ val,
               const T* __restrict__ B, int64_t ldb, MatSink<T> sink, int32_t chunk_rows) {
  using V = typename VecT<T, VEC>::type;
  extern __shared__ __align__(128) unsigned char rm_stage[];
  constexpr int GPB = BLOCK / G;  // row groups per CTA
  constexpr int U = NB >= 8 ? 1 : NB >= 4 ? 2 : 4;  // nnz per step (MLP = U*NB gathers per lane)
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const int gi = threadIdx.x / G;
  const int32_t r0 = tile_row[blockIdx.x], r1 = tile_row[blockIdx.x + 1];
  const int k = sink.k;
  const uint32_t rowbytes = (uint32_t)k * (uint32_t)sizeof(T);
  if (!BULK) chunk_rows = r1 - r0 > 0 ? r1 - r0 : 1;

 for (int32_t cr0 = r0; cr0 < r1; cr0 += chunk_rows) {  // block-uniform
  const int32_t cr1 = min(cr0 + chunk_rows, r1);
  for (int32_t rb = cr0; rb < cr1; rb += GPB) {  // block-uniform
    const int32_t r = rb + gi;
    const bool valid = r < cr1;
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
          } else if (BULK) {   // full vectors only (host: k % VEC == 0)
            const int64_t off = (int64_t)r * sink.ldc + c0;
            const V res = vaxpby(sink.alpha, acc[nb], sink.beta, *reinterpret_cast<const V*>(sink.C + off));
            *reinterpret_cast<V*>(sink.C + off) = res;
            *reinterpret_cast<V*>(rm_stage + (size_t)(r - cr0) * rowbytes + (size_t)c0 * sizeof(T)) = res;
          } else {
            emit_row_slice<T, VEC>(sink, (int64_t)r * sink.ldc + c0, c0, acc[nb]);
          }
        }
      }
    }
  }
  if (BULK) {
    fence_proxy_async_smem();   // my generic-proxy writes of the staged rows precede the bulk reads
    __syncthreads();
    // rows of the chunk that belong in the replicas: all but a HEAD / TAIL partial row (those travel as carries)
    int32_t a = cr0, b = cr1;
    if ((sink.flags & SBLAS_FLAG_HEAD_PARTIAL) && a == 0) ++a;
    if ((sink.flags & SBLAS_FLAG_TAIL_PARTIAL) && b == sink.rows) --b;
    if (sink.ldc == k) {      // the rows are one contiguous block of C
      if (threadIdx.x == 0 && b > a) {
        for (int d = 0; d < sink.out.n; ++d)
          bulk_s2g(sink.out.ptr[d] + (int64_t)a * sink.ldc, rm_stage + (size_t)(a - cr0) * rowbytes, (uint32_t)(b - a) * rowbytes);
        bulk_commit();
        bulk_wait_read_all();
      }
    } else if ((int)threadIdx.x < b - a) {   // pitched C: one bulk store per row and peer
      const int32_t row = a + (int)threadIdx.x;
      for (int d = 0; d < sink.out.n; ++d)
        bulk_s2g(sink.out.ptr[d] + (int64_t)row * sink.ldc, rm_stage + (size_t)(row - cr0) * rowbytes, rowbytes);
      bulk_commit();
      bulk_wait_read_all();
    }
    if (cr1 < r1) __syncthreads();   // the stage is free for the next chunk
  }
 }
  if (BULK) bulk_wait_all();   // the peers hold my rows before the CTA retires (the stream's next kernel signals them)
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
            emit_row_slice<T, VEC>(sink, row * sink.ldc + c0, c0, acc[r][nb]);
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
#define SBLAS_RB(NN, RR)                                                                          \
  spmm_rb_kernel<T, VEC, NN, RR, EXACT, BLOCK><<<grid, BLOCK, 0, st>>>(                            \
      rb.n_blockrows, rb.d_ptr, (const int2*)rb.d_colmask, (const T*)rb.d_val, B, ldb, sink)
  if (rb.R == 8) {  // NB <= 2 (the caller sizes the panels)
    if (NB <= 1) SBLAS_RB(1, 8);
    else SBLAS_RB(2, 8);
  } else {
    if (NB <= 1) SBLAS_RB(1, 4);
    else if (NB == 2) SBLAS_RB(2, 4);
    else SBLAS_RB(4, 4);
  }
#undef SBLAS_RB
  SBLAS_LAUNCH_CHECK();
  return 0;
}

// ---- ROWBLOCK, pipelined (persistent CTAs, producer warp + 8 consumer warps) ---------------------
// The kernel above is latency-bound: every warp walks blk_ptr -> {col,mask} -> B row -> old C as a
// chain of dependent global loads, and its instruction stream carries ~70 warp instructions per
// block for ~16 that are arithmetic.  Here a CTA walks CONSECUTIVE tiles of W block rows (so the
// B rows a tile gathered are still in L1 for the next one: neighbouring rows of banded / stencil
// inputs share most columns); warp 8 is a producer that keeps NSTAGE tiles of the A mirror in
// flight — the tile's {col,mask} and value ranges are contiguous (cp.async.bulk, SASS UBLKCP), its
// W+1 block pointers go by LDGSTS; one "full" mbarrier phase per stage collects all of it, one
// "empty" mbarrier per stage hands the stage back (no CTA-wide barrier in the loop, warps drift
// across stages).  Shared memory is kept small (2 x 12 KB stages + 16 KB) because what it takes
// is taken from L1, and the L1 hit rate of the B gather is what the kernel lives on.
// The 8 consumer warps are W x P: warp (w, p) owns block row w of the tile and column slice p
// (NB*32*VEC columns) of the panel, so k up to P*NB*32*VEC is covered by ONE pass over A.  A
// warp starts the copy of its R old C row slices (LDGSTS past L1, 16 bytes per lane) before the
// block loop and reads them back in the epilogue.  The B gather is covered by U blocks in flight
// per warp and 16-24 consumer warps per SM.  Instruction issue is the other bound, so the inner
// loop is lean: shared loads by 32-bit shared address (LDS.64 {col,mask}, LDS.128 values), one
// IMAD.WIDE for the B row address, unpredicated (column-clamped) LDG.128, a branch to an
// unpredicated body when all U masks are full, one predicated pass for the < U blocks left at
// the end of a block row: 24 instructions per block, 16 of them DMUL/DADD.
// (Tried and dropped: a per-warp cp.async ring for the gathered B rows — LDGSTS + LDS doubled the
// instruction count and the shared-memory wavefronts: 15.7 ms against 11.1 ms on the k=64 case;
// cyclic tile distribution: 9.8 ms against 8.9 ms; C tiles staged by TMA: 32 KB less L1.)
constexpr int kRbpConsumers = 8;                      // consumer warps per CTA
constexpr int kRbpBlock = (kRbpConsumers + 1) * 32;   // + the producer warp
constexpr int kRbpHeader = 128;                       // full[NSTAGE], empty[NSTAGE] barriers

template <typename T, int R>
__host__ __device__ inline size_t rbp_stage_bytes(int32_t cap_blocks) {
  return (size_t)cap_blocks * 8 + (size_t)cap_blocks * R * sizeof(T) + 64;
}
// per consumer warp: the old C values of its R rows (16 bytes per lane, row and slice)
template <int R>
__host__ __device__ constexpr int rbp_cbuf_bytes(int NB) { return R * NB * 512; }

// shared-memory loads by 32-bit shared address (volatile: never hoisted over the mbarrier waits)
__device__ __forceinline__ int2 lds_i2(uint32_t a) {
  int2 r;
  asm volatile("ld.shared.v2.s32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a));
  return r;
}
__device__ __forceinline__ int lds_i1(uint32_t a) {
  int r;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(r) : "r"(a));
  return r;
}
template <int R>
__device__ __forceinline__ void lds_vals(uint32_t a, double (&vv)[R]) {
#pragma unroll
  for (int r = 0; r < R; r += 2)
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(vv[r]), "=d"(vv[r + 1]) : "r"(a + r * 8));
}
template <int R>
__device__ __forceinline__ void lds_vals(uint32_t a, float (&vv)[R]) {
#pragma unroll
  for (int r = 0; r < R; r += 4)
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(vv[r]), "=f"(vv[r + 1]), "=f"(vv[r + 2]), "=f"(vv[r + 3]) : "r"(a + r * 4));
}
__device__ __forceinline__ double2 lds_v(uint32_t a, double2) {
  double2 r;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(a));
  return r;
}
__device__ __forceinline__ float4 lds_v(uint32_t a, float4) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(a));
  return r;
}

__device__ __forceinline__ void sts_v(uint32_t a, const double2& v) {
  asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(a), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ void sts_v(uint32_t a, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// shared (32-bit address) -> global bulk copy, see bulk_s2g
__device__ __forceinline__ void bulk_s2g_a(void* gdst, uint32_t ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}

// U blocks whose {col,mask} entries start at shared address a_cm and values at a_val (TAIL: only
// the first n < U exist): gathers first, then the multiply-adds; a branch takes an unpredicated
// body when all U masks are full.  bb[nb] = this lane's (clamped, always valid) column of slice nb
// in row 0 of B: lanes past the panel edge compute garbage that the epilogue never stores.
template <typename T, int VEC, int NB, int R, bool EXACT, int U, bool TAIL>
__device__ __forceinline__ void rbp_blocks(uint32_t a_cm, uint32_t a_val, int n, const char* const (&bb)[NB],
                                           uint32_t ldb_bytes, typename VecT<T, VEC>::type (&acc)[R][NB]) {
  using V = typename VecT<T, VEC>::type;
  constexpr int FULL = (1 << R) - 1;
  int mask[U];
  V bv[U][NB];
  int allmask = FULL;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (TAIL && u >= n) {
      mask[u] = 0;
    } else {
      const int2 cm = lds_i2(a_cm + u * 8);
      mask[u] = cm.y & kRbMaskBits;
      const uint64_t roff = (uint64_t)(uint32_t)cm.x * ldb_bytes;
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) bv[u][nb] = ldg_v(reinterpret_cast<const V*>(bb[nb] + roff));
    }
    allmask &= mask[u];
  }
  if (!TAIL && allmask == FULL) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      T vv[R];
      lds_vals<R>(a_val + u * (R * (int)sizeof(T)), vv);
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) vmac<EXACT>(vv[r], bv[u][nb], acc[r][nb]);
    }
  } else {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (mask[u] == 0) continue;
      T vv[R];
      lds_vals<R>(a_val + u * (R * (int)sizeof(T)), vv);
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (mask[u] & (1 << r)) {
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) vmac<EXACT>(vv[r], bv[u][nb], acc[r][nb]);
        }
    }
  }
}

constexpr int kRbpCPrefetchDist = 2;  // tiles ahead for the L2 prefetch of old C rows (SBLAS_FLAG_PREFETCH_C)

template <typename T, int VEC, int NB, int R, bool EXACT, int NSTAGE>
__global__ void __launch_bounds__(kRbpBlock, (R == 4 ? 4 : 3) - NB)
spmm_rbp_kernel(int64_t n_blockrows, const int32_t* __restrict__ blk_ptr, const int2* __restrict__ colmask,
                const T* __restrict__ bval, const T* __restrict__ B, uint32_t ldb_bytes, MatSink<T> sink,
                int32_t cap_blocks, int32_t kp, int32_t logP, int32_t prefetch, int32_t bulk) {
  // bulk: 0 = every lane stores its result slices itself (st.global to C and to every replica);
  //       1 = the results are parked in the warp's C buffer (the old C values they replace) and lane 0 pushes each
  //           512-byte row slice to every REPLICA with one TMA bulk store (the stitch of method 2 leaves the LSU);
  //       2 = ... and to the local C as well (no st.global left in the epilogue)
  using V = typename VecT<T, VEC>::type;
  constexpr int U = NB == 1 ? 4 : 2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* empty = full + NSTAGE;
  unsigned char* stages = smem_raw + kRbpHeader;
  const int W = kRbpConsumers >> logP;  // block rows per tile
  const size_t stage_bytes = rbp_stage_bytes<T, R>(cap_blocks);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&full[s], 32 + 1);             // producer lanes' LDGSTS groups + the expect_tx arrive
      mbar_init(&empty[s], kRbpConsumers);     // one arrive per consumer warp
    }
    mbar_fence_init();
  }
  __syncthreads();
  const int64_t n_tiles = (n_blockrows + W - 1) / W;
  // consecutive tiles per CTA: neighbouring block rows share most of their columns (banded /
  // stencil inputs), so the B rows of tile i are still in L1 when tile i+1 gathers them
  const int64_t per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int64_t tile0 = (int64_t)blockIdx.x * per_cta;
  const int my_tiles = (int)max((int64_t)0, min(per_cta, n_tiles - tile0));

  auto stage_ptrs = [&](int s, int2*& s_cm, T*& s_val, int32_t*& s_ptr) {
    unsigned char* p = stages + (size_t)s * stage_bytes;
    s_cm = reinterpret_cast<int2*>(p);
    p += (size_t)cap_blocks * 8;
    s_val = reinterpret_cast<T*>(p);
    p += (size_t)cap_blocks * R * sizeof(T);
    s_ptr = reinterpret_cast<int32_t*>(p);
  };

  if (warp == kRbpConsumers) {
    // ---- producer warp --------------------------------------------------------------------------
#pragma unroll 1
    for (int i = 0; i < my_tiles; ++i) {
      const int s = i % NSTAGE;
      const int64_t br0 = (tile0 + i) * W;
      const int32_t nbr = (int32_t)min((int64_t)W, n_blockrows - br0);
      const int32_t p0 = __ldg(blk_ptr + br0), p1 = __ldg(blk_ptr + br0 + nbr);
      if (i >= NSTAGE) mbar_wait(&empty[s], (uint32_t)(i / NSTAGE - 1) & 1u);
      int2* s_cm; T* s_val; int32_t* s_ptr;
      stage_ptrs(s, s_cm, s_val, s_ptr);
      const int32_t p0a = p0 & ~1;                 // {col,mask} entries are 8 bytes: bulk needs pairs
      const int32_t m0 = min((p0 + 1) & ~1, p1);   // aligned middle [m0, m1)
      const int32_t m1 = max(p1 & ~1, m0);
      if (lane == 0) {
        const uint32_t nb = (uint32_t)(p1 - p0);
        const uint32_t bytes = (uint32_t)(m1 - m0) * 8u + nb * (uint32_t)(R * sizeof(T));
        mbar_arrive_expect_tx(&full[s], bytes);
        if (m1 > m0) bulk_g2s(s_cm + (m0 - p0a), colmask + m0, (uint32_t)(m1 - m0) * 8u, &full[s]);
        if (nb) bulk_g2s(s_val + (size_t)(p0 - p0a) * R, bval + (size_t)p0 * R, nb * (uint32_t)(R * sizeof(T)), &full[s]);
      }
      if (lane == 0 && m0 > p0) cp_async8(s_cm + (p0 - p0a), colmask + p0);   // head (1 entry)
      if (lane == 1 && p1 > m1) cp_async8(s_cm + (m1 - p0a), colmask + m1);   // tail (1 entry)
      if (lane >= 2 && lane < 2 + nbr + 1) cp_async4(s_ptr + (lane - 2), blk_ptr + br0 + (lane - 2));
      cp_async_mbar_arrive_noinc(&full[s]);
      if (prefetch & 1) {
        // the B rows this tile will gather: pull them into L1 while the consumers are still on
        // the previous tile (first touches come from HBM, nobody else prefetches them).
        // (cp.async.bulk.prefetch.L2 per row instead: 12.5 ms against 9.3 ms without, TMA-bound.)
        mbar_wait(&full[s], (uint32_t)(i / NSTAGE) & 1u);
        const uint32_t rowb = (uint32_t)kp * (uint32_t)sizeof(T);
        for (int32_t j = (p0 - p0a) + lane; j < p1 - p0a; j += 32) {
          const char* brow = reinterpret_cast<const char*>(B) + (uint64_t)(uint32_t)s_cm[j].x * ldb_bytes;
          for (uint32_t o = 0; o < rowb; o += 128) prefetch_l1(brow + o);
        }
      }
    }
    return;
  }

  // ---- consumer warps: (w, p) = (block row of the tile, column slice of the panel) ---------------
  const int w = warp >> logP, p = warp & ((1 << logP) - 1);
  const int cbase = (p * NB * 32 + lane) * VEC;  // first column of this lane, slice nb adds nb*32*VEC
  bool colok[NB];
  const char* bb[NB];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int c0 = cbase + nb * 32 * VEC;
    colok[nb] = c0 < kp;
    bb[nb] = reinterpret_cast<const char*>(B + min(c0, kp - VEC));   // clamped: loads stay unpredicated
  }
  const uint32_t stages_a = smem_u32(stages);
  const uint32_t off_val = (uint32_t)cap_blocks * 8u, off_ptr = off_val + (uint32_t)cap_blocks * (uint32_t)(R * sizeof(T));
  // this warp's C buffer, behind the stages: lane l keeps the 16 bytes it copied at (r*NB+nb)*512 + l*16
  const uint32_t a_C = stages_a + (uint32_t)(NSTAGE * stage_bytes) + (uint32_t)(warp * rbp_cbuf_bytes<R>(NB) + lane * 16);

#pragma unroll 1
  for (int i = 0; i < my_tiles; ++i) {
    const int s = i % NSTAGE;
    const uint32_t st_a = stages_a + (uint32_t)s * (uint32_t)stage_bytes;
    mbar_wait(&full[s], (uint32_t)(i / NSTAGE) & 1u);
    const int64_t br = (tile0 + i) * W + w;
    if ((prefetch & 2) && i + kRbpCPrefetchDist < my_tiles) {
      // the old C rows this warp will need kRbpCPrefetchDist tiles from now: into L2 today, so that the
      // LDGSTS at the top of that tile returns in an L2 round trip and not a DRAM one (the block loop of one
      // tile is shorter than the loaded DRAM latency, so the C fetch was exposed at every epilogue)
      const int64_t brn = br + (int64_t)kRbpCPrefetchDist * W;
      if (brn < n_blockrows) {
        const int lines = (NB * 32 * VEC * (int)sizeof(T) + 127) / 128;   // 128-byte lines per row slice
        for (int idx = lane; idx < R * lines; idx += 32) {
          const int r = idx / lines, ln = idx - r * lines;
          const int64_t row = brn * R + r;
          const int c0 = p * NB * 32 * VEC + ln * (128 / (int)sizeof(T));
          if (row < sink.rows && c0 < kp) prefetch_l2(sink.C + row * sink.ldc + c0);
        }
      }
    }
    if (br < n_blockrows) {
      const int32_t p0a = lds_i1(st_a + off_ptr) & ~1;
      const int2 bp = make_int2(lds_i1(st_a + off_ptr + w * 4), lds_i1(st_a + off_ptr + w * 4 + 4));
      if (bulk) {   // the bulk stores of the previous tile have read the C buffer before it is refilled
        if (lane == 0) bulk_wait_read_all();
        __syncwarp();
      }
      // old C values of my R rows: LDGSTS past L1 now, read back in the epilogue
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int64_t row = br * R + r;
        if (row < sink.rows) {
#pragma unroll
          for (int nb = 0; nb < NB; ++nb)
            if (colok[nb]) cp_async16_cg(a_C + (uint32_t)((r * NB + nb) * 512), sink.C + row * sink.ldc + cbase + nb * 32 * VEC);
        }
      }
      cp_async_commit();
      V acc[R][NB];
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) vzero(acc[r][nb]);
      uint32_t a_cm = st_a + (uint32_t)(bp.x - p0a) * 8u;
      uint32_t a_val = st_a + off_val + (uint32_t)(bp.x - p0a) * (uint32_t)(R * sizeof(T));
      const uint32_t a_end = st_a + (uint32_t)(bp.y - p0a) * 8u;
#pragma unroll 1
      for (; a_cm + U * 8 <= a_end; a_cm += U * 8, a_val += U * R * sizeof(T))
        rbp_blocks<T, VEC, NB, R, EXACT, U, false>(a_cm, a_val, U, bb, ldb_bytes, acc);
      if (a_cm < a_end)
        rbp_blocks<T, VEC, NB, R, EXACT, U, true>(a_cm, a_val, (int)(a_end - a_cm) >> 3, bb, ldb_bytes, acc);
      cp_async_wait<0>();
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int64_t row = br * R + r;
        if (row >= sink.rows) break;
        const int slot = sink.partial_slot((int32_t)row);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const int c0 = cbase + nb * 32 * VEC;
          if (colok[nb]) {
            if (slot >= 0) {
              *reinterpret_cast<V*>(sink.carry + (int64_t)slot * sink.kpad + c0) = acc[r][nb];
            } else {
              const int64_t off = row * sink.ldc + c0;
              const V cv = lds_v(a_C + (uint32_t)((r * NB + nb) * 512), V());
              const V res = vaxpby(sink.alpha, acc[r][nb], sink.beta, cv);
              if (bulk) {
                sts_v(a_C + (uint32_t)((r * NB + nb) * 512), res);
                if (bulk == 1) *reinterpret_cast<V*>(sink.C + off) = res;
              } else {
                *reinterpret_cast<V*>(sink.C + off) = res;
                for (int d = 0; d < sink.out.n; ++d) *reinterpret_cast<V*>(sink.out.ptr[d] + off) = res;
              }
            }
          }
        }
      }
      if (bulk) {
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          const int wc0 = p * NB * 32 * VEC;                 // first column of the warp's slices
          const uint32_t a_W = a_C - (uint32_t)(lane * 16);
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const int64_t row = br * R + r;
            if (row >= sink.rows) break;
            if (sink.partial_slot((int32_t)row) >= 0) continue;
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
              const int c0 = wc0 + nb * 32 * VEC;
              if (c0 >= kp) break;
              const uint32_t bytes = (uint32_t)min(32 * VEC, kp - c0) * (uint32_t)sizeof(T);
              const int64_t off = row * sink.ldc + c0;
              const uint32_t src = a_W + (uint32_t)((r * NB + nb) * 512);
              if (bulk == 2) bulk_s2g_a(sink.C + off, src, bytes);
              for (int d = 0; d < sink.out.n; ++d) bulk_s2g_a(sink.out.ptr[d] + off, src, bytes);
            }
          }
          bulk_commit();
        }
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }
  if (bulk && lane == 0) bulk_wait_all();   // every row slice has landed (locally and on the peers) before the warp retires
}

// P = column slices per CTA (1, 2, 4); panel = P*NB*32*VEC columns per pass over A
template <typename T, int VEC, bool EXACT, int R>
static int launch_rbp_r(cudaStream_t st, const sblasPlan_st* plan, int NB, int logP, int stages, const T* B,
                        int64_t ldb, const MatSink<T>& sink, int32_t kp) {
  const RowBlocks& rb = plan->rb;
  const int W = kRbpConsumers >> logP;
  const int32_t tile_blocks = logP == 0 ? rb.max_tile_blocks : logP == 1 ? rb.max_tile_blocks4 : rb.max_tile_blocks2;
  const int32_t cap_blocks = (tile_blocks + 2 + 1) & ~1;
  auto smem_for = [&](int n) {
    return kRbpHeader + n * rbp_stage_bytes<T, R>(cap_blocks) + (size_t)kRbpConsumers * rbp_cbuf_bytes<R>(NB <= 1 ? 1 : 2);
  };
  // stages of the A mirror in flight (2..4; R = 8 mirrors: 2), while 2 CTAs still fit an SM
  int ns = R == 4 ? std::min(std::max(stages, 2), 4) : 2;
  while (ns > 2 && smem_for(ns) > 100 * 1024) --ns;
  const void* kern;
#define SBLAS_RBP(NS) (NB <= 1 ? (const void*)spmm_rbp_kernel<T, VEC, 1, R, EXACT, NS> \
                               : (const void*)spmm_rbp_kernel<T, VEC, 2, R, EXACT, NS>)
  if (ns == 2) kern = SBLAS_RBP(2);
  else if (ns == 3) kern = SBLAS_RBP((R == 4 ? 3 : 2));
  else kern = SBLAS_RBP((R == 4 ? 4 : 2));
#undef SBLAS_RBP
  const size_t smem = smem_for(ns);
  SBLAS_CHECK(smem <= 200 * 1024, "spmm rowblock pipe: tile needs %zu bytes of shared memory", smem);
  // experiment hooks: SBLAS_RBP_CARVEOUT = percent of the 256 KB L1/shared array given to shared memory,
  // SBLAS_RBP_PER_SM = resident CTAs per SM the persistent grid is sized for
  static const int env_carve = [] { const char* e = getenv("SBLAS_RBP_CARVEOUT"); return e ? atoi(e) : -1; }();
  static const int env_per_sm = [] { const char* e = getenv("SBLAS_RBP_PER_SM"); return e ? atoi(e) : 0; }();
  if (env_carve >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, env_carve);
  int per_sm = 0, n_sm = kNumSM;
  if (int rr = kernel_residency(kern, kRbpBlock, smem, &per_sm, &n_sm)) return rr;
  if (env_per_sm > 0) per_sm = std::min(per_sm, env_per_sm);
  const int64_t n_tiles = ceil_div(rb.n_blockrows, W);
  const int grid = (int)std::min<int64_t>(n_tiles, (int64_t)n_sm * per_sm);
  const int2* cm = (const int2*)rb.d_colmask;
  const T* bv = (const T*)rb.d_val;
  const uint32_t ldb_bytes = (uint32_t)(ldb * (int64_t)sizeof(T));
  const int32_t prefetch = ((sink.flags & SBLAS_FLAG_NO_PREFETCH_B) ? 0
                            : ((sink.flags & SBLAS_FLAG_PREFETCH_B) || (NB <= 1 && logP == 0)) ? 1 : 0) |
                           ((sink.flags & SBLAS_FLAG_PREFETCH_C) ? 2 : 0);
  // epilogue stores: per-lane st.global by default.  TMA bulk stores of the 512-byte row slices (SBLAS_RBP_BULK=1:
  // to the replicas, =2: to the local C as well; needs a 16-byte row pitch) are parity-tested but measured level or
  // slower: b32m 1 GPU 8.44 -> 8.72 ms (=2), 2 GPUs 7.05 -> 7.16 ms, 4 GPUs 9.70 -> 9.77 ms (=1)
  // (profiles/r02_exp_b32m_rbp_bulk_n1.log, r02_bench_b32m_n{2,4}*.json): a 512-byte slice is too small a TMA
  // transaction to beat 32 coalesced 16-byte stores, unlike the 16-32 KB chunks of the CSR kernel's stitch.
  static const int env_bulk = [] { const char* e = getenv("SBLAS_RBP_BULK"); return e ? atoi(e) : 0; }();
  int32_t bulk = env_bulk > 0 ? env_bulk : ((sink.flags & SBLAS_FLAG_STITCH_BULK) && sink.out.n > 0 ? 1 : 0);
  if ((sink.flags & SBLAS_FLAG_STITCH_DIRECT) || (sink.ldc * (int64_t)sizeof(T)) % 16 != 0) bulk = 0;
  for (int d = 0; d < sink.out.n; ++d)
    if (!aligned16(sink.out.ptr[d])) bulk = 0;
  void* args[] = {(void*)&rb.n_blockrows, (void*)&rb.d_ptr, (void*)&cm, (void*)&bv, (void*)&B, (void*)&ldb_bytes,
                  (void*)&sink, (void*)&cap_blocks, (void*)&kp, (void*)&logP, (void*)&prefetch, (void*)&bulk};
  SBLAS_CUDA(cudaLaunchKernel(kern, dim3(grid), dim3(kRbpBlock), args, smem, st));
  ++g_launches;
  return 0;
}

// shared memory the pipelined kernel needs for this plan (2 stages); it falls back to the
// unpiped kernel when a tile of the mirror does not fit
template <typename T>
static size_t rbp_smem_bytes(const sblasPlan_st* plan, int NB, int logP) {
  const RowBlocks& rb = plan->rb;
  const int32_t tile_blocks = logP == 0 ? rb.max_tile_blocks : logP == 1 ? rb.max_tile_blocks4 : rb.max_tile_blocks2;
  const int32_t cap_blocks = (tile_blocks + 2 + 1) & ~1;
  const size_t stage = rb.R == 8 ? rbp_stage_bytes<T, 8>(cap_blocks) : rbp_stage_bytes<T, 4>(cap_blocks);
  const size_t cbuf = rb.R == 8 ? rbp_cbuf_bytes<8>(NB <= 1 ? 1 : 2) : rbp_cbuf_bytes<4>(NB <= 1 ? 1 : 2);
  return kRbpHeader + 2 * stage + (size_t)kRbpConsumers * cbuf;
}

template <typename T, int VEC, bool EXACT>
static int launch_rbp(cudaStream_t st, const sblasPlan_st* plan, int NB, int logP, int stages, const T* B,
                      int64_t ldb, const MatSink<T>& sink, int32_t kp) {
  return plan->rb.R == 8 ? launch_rbp_r<T, VEC, EXACT, 8>(st, plan, NB, logP, stages, B, ldb, sink, kp)
                         : launch_rbp_r<T, VEC, EXACT, 4>(st, plan, NB, logP, stages, B, ldb, sink, kp);
}

// ---- ROWBLOCK, B rows staged per tile (TILEB) -----------------------------------------------------
// ncu on the kernel above (k = 64, banded): issue slots 55 %, FP64 46 %, LSU wavefronts 61 %, and the
// rest of the time warps sit in long_scoreboard on the B gathers that miss L1 — every warp fetches
// "its" B row slices itself although the 8 block rows of a tile share most columns (280 gathers
// for 63 distinct rows on the 32-wide band).  Here the mirror also carries, per tile of 8 block
// rows, the ascending list of its DISTINCT columns, and every block the position of its column in
// that list (spmm_rb.cu, rows with ascending columns only).  The producer warp stages the tile's A
// blocks AND those B row slices (one cp.async.bulk per distinct column, kp*sizeof(T) bytes) behind
// the stage's mbarrier, two tiles ahead; the consumer warps take B from shared memory (LDS.128,
// conflict-free, no global latency in the block loop at all).  Used when the tile's column list
// fits two stages in ~100 KB (narrow bands, k of one or two slices); else the kernel above.
template <typename T, int R>
__host__ __device__ inline size_t rbt_stage_bytes(int32_t cap_blocks, int32_t cap_cols, int NB) {
  return rbp_stage_bytes<T, R>(cap_blocks) + (size_t)cap_cols * NB * 512;
}

// tag = colmask.y = presence mask | (position of the block's column in the tile list) << 9, so
// the byte offset of the staged B row slice is (tag & ~0x1ff) * NB
template <typename T, int VEC, int NB, int R, bool EXACT, int U, bool TAIL>
__device__ __forceinline__ void rbt_blocks(uint32_t a_cm, uint32_t a_val, int n, uint32_t a_B,
                                           typename VecT<T, VEC>::type (&acc)[R][NB]) {
  using V = typename VecT<T, VEC>::type;
  constexpr int FULL = (1 << R) - 1;
  int tag[U];
  int allmask = FULL;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    tag[u] = (TAIL && u >= n) ? 0 : lds_i1(a_cm + u * 8 + 4);
    allmask &= tag[u];
  }
  if (!TAIL && allmask == FULL) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t a_row = a_B + ((uint32_t)tag[u] & ~0x1ffu) * NB;
      T vv[R];
      lds_vals<R>(a_val + u * (R * (int)sizeof(T)), vv);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const V bv = lds_v(a_row + nb * 512, V());
#pragma unroll
        for (int r = 0; r < R; ++r) vmac<EXACT>(vv[r], bv, acc[r][nb]);
      }
    }
  } else {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if ((tag[u] & kRbMaskBits) == 0) continue;
      const uint32_t a_row = a_B + ((uint32_t)tag[u] & ~0x1ffu) * NB;
      T vv[R];
      lds_vals<R>(a_val + u * (R * (int)sizeof(T)), vv);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const V bv = lds_v(a_row + nb * 512, V());
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (tag[u] & (1 << r)) vmac<EXACT>(vv[r], bv, acc[r][nb]);
      }
    }
  }
}

template <typename T, int VEC, int NB, int R, bool EXACT>
__global__ void __launch_bounds__(kRbpBlock, 2)
spmm_rbt_kernel(int64_t n_blockrows, const int32_t* __restrict__ blk_ptr, const int2* __restrict__ colmask,
                const T* __restrict__ bval, const int32_t* __restrict__ tile_colptr,
                const int32_t* __restrict__ tile_cols, const T* __restrict__ B, uint32_t ldb_bytes,
                MatSink<T> sink, int32_t cap_blocks, int32_t cap_cols, int32_t kp) {
  using V = typename VecT<T, VEC>::type;
  constexpr int NSTAGE = 2, U = 4, W = kRbpConsumers, ROWB = NB * 512;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* empty = full + NSTAGE;
  unsigned char* stages = smem_raw + kRbpHeader;
  const size_t stage_bytes = rbt_stage_bytes<T, R>(cap_blocks, cap_cols, NB);
  const uint32_t off_val = (uint32_t)cap_blocks * 8u, off_ptr = off_val + (uint32_t)cap_blocks * (uint32_t)(R * sizeof(T));
  const uint32_t off_B = off_ptr + 64u;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&full[s], 32 + 1);
      mbar_init(&empty[s], kRbpConsumers);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const int64_t n_tiles = (n_blockrows + W - 1) / W;
  const int64_t per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int64_t tile0 = (int64_t)blockIdx.x * per_cta;
  const int my_tiles = (int)max((int64_t)0, min(per_cta, n_tiles - tile0));

  if (warp == kRbpConsumers) {
    // ---- producer warp --------------------------------------------------------------------------
    const uint32_t rowb = (uint32_t)kp * (uint32_t)sizeof(T);
#pragma unroll 1
    for (int i = 0; i < my_tiles; ++i) {
      const int s = i % NSTAGE;
      const int64_t t = tile0 + i;
      const int64_t br0 = t * W;
      const int32_t nbr = (int32_t)min((int64_t)W, n_blockrows - br0);
      const int32_t p0 = __ldg(blk_ptr + br0), p1 = __ldg(blk_ptr + br0 + nbr);
      const int32_t c0 = __ldg(tile_colptr + t), c1 = __ldg(tile_colptr + t + 1);
      if (i >= NSTAGE) mbar_wait(&empty[s], (uint32_t)(i / NSTAGE - 1) & 1u);
      unsigned char* st = stages + (size_t)s * stage_bytes;
      int2* s_cm = reinterpret_cast<int2*>(st);
      T* s_val = reinterpret_cast<T*>(st + off_val);
      int32_t* s_ptr = reinterpret_cast<int32_t*>(st + off_ptr);
      unsigned char* s_B = st + off_B;
      const int32_t p0a = p0 & ~1;
      const int32_t m0 = min((p0 + 1) & ~1, p1);
      const int32_t m1 = max(p1 & ~1, m0);
      if (lane == 0) {
        const uint32_t nb = (uint32_t)(p1 - p0);
        const uint32_t bytes = (uint32_t)(m1 - m0) * 8u + nb * (uint32_t)(R * sizeof(T)) + (uint32_t)(c1 - c0) * rowb;
        mbar_arrive_expect_tx(&full[s], bytes);
        if (m1 > m0) bulk_g2s(s_cm + (m0 - p0a), colmask + m0, (uint32_t)(m1 - m0) * 8u, &full[s]);
        if (nb) bulk_g2s(s_val + (size_t)(p0 - p0a) * R, bval + (size_t)p0 * R, nb * (uint32_t)(R * sizeof(T)), &full[s]);
      }
      __syncwarp();
      // the tile's distinct B rows, one bulk copy each (after the expect_tx above)
      for (int32_t j = c0 + lane; j < c1; j += 32) {
        const int32_t col = __ldg(tile_cols + j);
        bulk_g2s(s_B + (size_t)(j - c0) * ROWB, reinterpret_cast<const char*>(B) + (uint64_t)(uint32_t)col * ldb_bytes,
                 rowb, &full[s]);
      }
      if (lane == 0 && m0 > p0) cp_async8(s_cm + (p0 - p0a), colmask + p0);
      if (lane == 1 && p1 > m1) cp_async8(s_cm + (m1 - p0a), colmask + m1);
      if (lane >= 2 && lane < 2 + nbr + 1) cp_async4(s_ptr + (lane - 2), blk_ptr + br0 + (lane - 2));
      cp_async_mbar_arrive_noinc(&full[s]);
    }
    return;
  }

  // ---- consumer warps: warp w owns block row w of the tile ----------------------------------------
  const int w = warp;
  const int cbase = lane * VEC;
  bool colok[NB];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) colok[nb] = cbase + nb * 32 * VEC < kp;
  const uint32_t stages_a = smem_u32(stages);
  const uint32_t a_C = stages_a + (uint32_t)(NSTAGE * stage_bytes) + (uint32_t)(warp * rbp_cbuf_bytes<R>(NB) + lane * 16);

#pragma unroll 1
  for (int i = 0; i < my_tiles; ++i) {
    const int s = i % NSTAGE;
    const uint32_t st_a = stages_a + (uint32_t)s * (uint32_t)stage_bytes;
    const int64_t br = (tile0 + i) * W + w;
    if (br < n_blockrows) {  // old C values first: they do not depend on the stage
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int64_t row = br * R + r;
        if (row < sink.rows) {
#pragma unroll
          for (int nb = 0; nb < NB; ++nb)
            if (colok[nb]) cp_async16_cg(a_C + (uint32_t)((r * NB + nb) * 512), sink.C + row * sink.ldc + cbase + nb * 32 * VEC);
        }
      }
      cp_async_commit();
    }
    mbar_wait(&full[s], (uint32_t)(i / NSTAGE) & 1u);
    if (br < n_blockrows) {
      const int32_t p0a = lds_i1(st_a + off_ptr) & ~1;
      const int2 bp = make_int2(lds_i1(st_a + off_ptr + w * 4), lds_i1(st_a + off_ptr + w * 4 + 4));
      V acc[R][NB];
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) vzero(acc[r][nb]);
      uint32_t a_cm = st_a + (uint32_t)(bp.x - p0a) * 8u;
      uint32_t a_val = st_a + off_val + (uint32_t)(bp.x - p0a) * (uint32_t)(R * sizeof(T));
      const uint32_t a_end = st_a + (uint32_t)(bp.y - p0a) * 8u;
      const uint32_t a_B = st_a + off_B + (uint32_t)lane * 16u;
#pragma unroll 1
      for (; a_cm + U * 8 <= a_end; a_cm += U * 8, a_val += U * R * sizeof(T))
        rbt_blocks<T, VEC, NB, R, EXACT, U, false>(a_cm, a_val, U, a_B, acc);
      if (a_cm < a_end)
        rbt_blocks<T, VEC, NB, R, EXACT, U, true>(a_cm, a_val, (int)(a_end - a_cm) >> 3, a_B, acc);
      cp_async_wait<0>();
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int64_t row = br * R + r;
        if (row >= sink.rows) break;
        const int slot = sink.partial_slot((int32_t)row);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const int c0 = cbase + nb * 32 * VEC;
          if (colok[nb]) {
            if (slot >= 0) {
              *reinterpret_cast<V*>(sink.carry + (int64_t)slot * sink.kpad + c0) = acc[r][nb];
            } else {
              const int64_t off = row * sink.ldc + c0;
              const V cv = lds_v(a_C + (uint32_t)((r * NB + nb) * 512), V());
              const V res = vaxpby(sink.alpha, acc[r][nb], sink.beta, cv);
              *reinterpret_cast<V*>(sink.C + off) = res;
              for (int d = 0; d < sink.out.n; ++d) *reinterpret_cast<V*>(sink.out.ptr[d] + off) = res;
            }
          }
        }
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }
}

// shared memory of the TILEB kernel for this plan (0: no tile column lists)
template <typename T>
static size_t rbt_smem_bytes(const sblasPlan_st* plan, int NB) {
  const RowBlocks& rb = plan->rb;
  if (rb.max_tile_cols <= 0 || rb.R != 4) return 0;
  const int32_t cap_blocks = (rb.max_tile_blocks + 2 + 1) & ~1;
  return kRbpHeader + 2 * rbt_stage_bytes<T, 4>(cap_blocks, rb.max_tile_cols, NB) +
         (size_t)kRbpConsumers * rbp_cbuf_bytes<4>(NB);
}

template <typename T, int VEC, bool EXACT>
static int launch_rbt(cudaStream_t st, const sblasPlan_st* plan, int NB, const T* B, int64_t ldb,
                      const MatSink<T>& sink, int32_t kp) {
  const RowBlocks& rb = plan->rb;
  constexpr int R = 4;
  NB = NB <= 1 ? 1 : 2;
  const int32_t cap_blocks = (rb.max_tile_blocks + 2 + 1) & ~1;
  const int32_t cap_cols = rb.max_tile_cols;
  const void* kern = NB == 1 ? (const void*)spmm_rbt_kernel<T, VEC, 1, R, EXACT>
                             : (const void*)spmm_rbt_kernel<T, VEC, 2, R, EXACT>;
  const size_t smem = rbt_smem_bytes<T>(plan, NB);
  SBLAS_CHECK(smem > 0 && smem <= 220 * 1024, "spmm rowblock tileb: tile needs %zu bytes of shared memory", smem);
  int per_sm = 0, n_sm = kNumSM;
  if (int rr = kernel_residency(kern, kRbpBlock, smem, &per_sm, &n_sm)) return rr;
  const int64_t n_tiles = ceil_div(rb.n_blockrows, kRbpConsumers);
  const int grid = (int)std::min<int64_t>(n_tiles, (int64_t)n_sm * per_sm);
  const int2* cm = (const int2*)rb.d_colmask;
  const T* bv = (const T*)rb.d_val;
  const uint32_t ldb_bytes = (uint32_t)(ldb * (int64_t)sizeof(T));
  void* args[] = {(void*)&rb.n_blockrows, (void*)&rb.d_ptr, (void*)&cm, (void*)&bv, (void*)&rb.d_tile_colptr,
                  (void*)&rb.d_tile_cols, (void*)&B, (void*)&ldb_bytes, (void*)&sink, (void*)&cap_blocks,
                  (void*)&cap_cols, (void*)&kp};
  SBLAS_CUDA(cudaLaunchKernel(kern, dim3(grid), dim3(kRbpBlock), args, smem, st));
  ++g_launches;
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
  extern __shared__ __align__(128) unsigned char smem_raw[];
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
  // the stitch by TMA bulk stores: wide rows (one warp per row), whole vectors, 16-byte row pitch; the flag
  // SBLAS_FLAG_STITCH_DIRECT (multicast replicas, comparisons) and SBLAS_STITCH_BULK=0 keep the per-lane stores
  static const int env_bulk = [] { const char* e = getenv("SBLAS_STITCH_BULK"); return e ? atoi(e) : 1; }();
  if (VEC > 1 && G == 32 && sink.out.n > 0 && env_bulk && !(sink.flags & SBLAS_FLAG_STITCH_DIRECT) &&
      sink.k % VEC == 0 && (sink.ldc * (int64_t)sizeof(T)) % 16 == 0) {
    constexpr int GPB = kSpmmBlock / 32;
    const size_t rowbytes = (size_t)sink.k * sizeof(T);
    static const int env_kb = [] { const char* e = getenv("SBLAS_STITCH_CHUNK_KB"); const int v = e ? atoi(e) : 0; return v > 0 ? std::min(v, 160) : 32; }();
    int32_t chunk = (int32_t)(((size_t)env_kb * 1024) / rowbytes) / GPB * GPB;
    chunk = std::max(chunk, GPB);
    const size_t smem = (size_t)chunk * rowbytes;
#define SBLAS_RMB(NN)                                                                                       \
    {                                                                                                       \
      auto kern = spmm_rm_kernel<T, VEC, 32, NN, EXACT, kSpmmBlock, true>;                                  \
      if (smem > 48 * 1024) SBLAS_CUDA(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
      kern<<<tt->n_tiles, kSpmmBlock, smem, st>>>(tt->d_tile_row, rp, ci, v, B, ldb, sink, chunk);          \
    }
    if (NB == 1) SBLAS_RMB(1)
    else if (NB == 2) SBLAS_RMB(2)
    else if (NB <= 4) SBLAS_RMB(4)
    else SBLAS_RMB(8)
#undef SBLAS_RMB
    SBLAS_LAUNCH_CHECK();
    return 0;
  }
#define SBLAS_RM(GG, NN)                                                                         \
  spmm_rm_kernel<T, VEC, GG, NN, EXACT, kSpmmBlock>                                              \
      <<<tt->n_tiles, kSpmmBlock, 0, st>>>(tt->d_tile_row, rp, ci, v, B, ldb, sink, 0)
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

// process-wide default arithmetic of SpMM: -1 = not decided yet (environment), 0 = FMA, 1 = exact
static int g_spmm_exact = -1;
static int spmm_exact_default() {
  if (g_spmm_exact < 0) {
    const char* e = getenv("SBLAS_EXACT");
    g_spmm_exact = (e && atoi(e) != 0) ? 1 : 0;
  }
  return g_spmm_exact;
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

  EphemeralPlan eph;  // released on every exit path
  eph.st = st;
  if (!plan) {
    if (sblas_plan_create_i32(st, rows, nnz, rp, &eph.p)) return 4;
    plan = eph.p;
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
  // arithmetic: fused multiply-add by default (the north star grants fp64 <= 1e-12 / fp32 <= 1e-5 per row against
  // |A||x| for reduction-order differences, and the FP64 pipe issues one DFMA where the exact mode needs DMUL +
  // DADD: k = 256 banded 7.4 -> 6.2 ms); SBLAS_FLAG_EXACT (or SBLAS_EXACT=1 in the environment, for callers that
  // go through the C++ headers, or
  // sblas_set_spmm_exact(1)) keeps the separately rounded mul/add that is bit-identical to spmm.h:56-68
  const bool exact = (flags & SBLAS_FLAG_EXACT) ? true : (flags & SBLAS_FLAG_FMA) ? false : spmm_exact_default() != 0;

  if (rc == 0 && layout == SBLAS_ROW_MAJOR) {
    constexpr int VF = Vec16<T>::N;
    bool vec = aligned16(B) && aligned16(C) && (ldb % VF == 0) && (ldc % VF == 0) &&
               ldb >= ((k + VF - 1) / VF) * VF && ldc >= ((k + VF - 1) / VF) * VF &&
               (sink.carry == nullptr || aligned16(sink.carry));
    for (int d = 0; d < sink.out.n; ++d) vec = vec && aligned16(sink.out.ptr[d]);
    const int VEC = vec ? VF : 1;
    // row-block mirror (R x 1 BCSR, built by sblas_plan_build_rowblocks_*): one B load feeds up to
    // R adjacent rows.  Taken when the plan holds it for this value type, the operands are
    // vectorisable and a row slice keeps at least half a warp busy (k >= 16*VEC); AUTO or SBLAS_SPMM_ROWBLOCK.
    const int want = o ? o->variant : SBLAS_VARIANT_AUTO;
    const bool rb_ok = vec && plan->rb.n_blocks > 0 && plan->rb.is_f64 == (sizeof(T) == 8) && k >= 16 * VF;
    if ((want == SBLAS_SPMM_ROWBLOCK || want == SBLAS_SPMM_ROWBLOCK_PIPE || want == SBLAS_SPMM_ROWBLOCK_TILEB) && !rb_ok) {
      set_error("spmm: SBLAS_SPMM_ROWBLOCK needs a plan with row blocks of this value type, 16-byte aligned operands and k >= %d", 16 * VF);
      rc = 2;
    }
    if (rc == 0 && rb_ok && (want == SBLAS_VARIANT_AUTO || want == SBLAS_SPMM_ROWBLOCK || want == SBLAS_SPMM_ROWBLOCK_PIPE ||
                             want == SBLAS_SPMM_ROWBLOCK_TILEB)) {
      // pipelined flavour: needs the tile statistics and a B pitch below 4 GB.  A pass over A covers
      // P*NB slices of 32*VF columns.  AUTO: P = 1 (tiles of 8 block rows keep the most B rows
      // in L1) and NB = 2 as soon as there are two slices, i.e. k = 256 fp64 takes two passes
      // (7.3 ms on config 5 against 8.4 / 8.7 ms for one pass with P = 2 / 4); opts->lanes =
      // 1/2/4 forces P.  SBLAS_SPMM_ROWBLOCK keeps the unpiped kernel.
      bool piped = want != SBLAS_SPMM_ROWBLOCK && plan->rb.max_tile_blocks > 0 &&
                   ldb * (int64_t)sizeof(T) < (1ll << 32);
      const int slice = 32 * VF;
      int NBp = 1, logP = 0;
      if (piped) {
        const int n_slices = (k + slice - 1) / slice;
        if (o && (o->lanes == 1 || o->lanes == 2 || o->lanes == 4)) logP = o->lanes == 1 ? 0 : o->lanes == 2 ? 1 : 2;
        NBp = n_slices > (1 << logP) ? 2 : 1;
      }
      if (piped && rbp_smem_bytes<T>(plan, NBp, logP) > 100 * 1024) piped = false;  // >= 2 CTAs per SM or not at all
      // B rows staged once per tile: only on request (SBLAS_SPMM_ROWBLOCK_TILEB).  Measured on b32m
      // it is level with the pipe kernel (8.6-9.0 ms against 8.9 ms) although its block loop has no
      // global load left, so AUTO keeps the kernel with the smaller shared-memory footprint.
      const size_t tb = (piped && logP == 0 && want == SBLAS_SPMM_ROWBLOCK_TILEB) ? rbt_smem_bytes<T>(plan, NBp) : 0;
      const bool tileb = tb > 0 && tb <= 220u * 1024u;
      if (want == SBLAS_SPMM_ROWBLOCK_TILEB && !tileb) {
        set_error("spmm: SBLAS_SPMM_ROWBLOCK_TILEB needs tile column lists (rows with ascending columns, R = 4) that fit shared memory");
        rc = 2;
      }
      const int rbpanel = piped ? slice * NBp * (1 << logP) : slice * (plan->rb.R == 8 ? 2 : 4);
      for (int32_t c_off = 0; c_off < k && rc == 0; c_off += rbpanel) {
        const int kp = std::min<int32_t>(rbpanel, k - c_off);
        const int NB = (kp + 32 * VF - 1) / (32 * VF);
        MatSink<T> ps = sink;
        ps.C = C + c_off;
        ps.k = kp;
        for (int d = 0; d < ps.out.n; ++d) ps.out.ptr[d] = sink.out.ptr[d] + c_off;
        if (ps.carry) ps.carry = sink.carry + c_off;
        if (tileb && kp % VF == 0)
          rc = exact ? launch_rbt<T, VF, true>(st, plan, NBp, B + c_off, ldb, ps, kp)
                     : launch_rbt<T, VF, false>(st, plan, NBp, B + c_off, ldb, ps, kp);
        else if (piped && kp % VF == 0)
          rc = exact ? launch_rbp<T, VF, true>(st, plan, NBp, logP, o ? o->stages : 0, B + c_off, ldb, ps, kp)
                     : launch_rbp<T, VF, false>(st, plan, NBp, logP, o ? o->stages : 0, B + c_off, ldb, ps, kp);
        else
          rc = exact ? launch_rb<T, VF, true>(st, plan, std::min(NB, 4), B + c_off, ldb, ps)
                     : launch_rb<T, VF, false>(st, plan, std::min(NB, 4), B + c_off, ldb, ps);
      }
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
  return rc;
}

}  // namespace sblas

using namespace sblas;

extern "C" {
int sblas_set_spmm_exact(int on) {
  const int prev = spmm_exact_default();
  g_spmm_exact = on ? 1 : 0;
  return prev;
}
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
