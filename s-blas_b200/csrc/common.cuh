// common.cuh — shared device/host helpers for libsblas_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "sblas_b200.h"

namespace sblas {

// ---- error plumbing -------------------------------------------------------------------
// The reference prints and exit(-1)s on any CUDA failure (utility.h:27-83).  The C ABI
// returns a code and keeps the message; the C++ drop-in headers turn it back into
// print-and-exit.
void set_error(const char* fmt, ...);
const char* get_error();

#define SBLAS_CUDA(call)                                                                   \
  do {                                                                                     \
    cudaError_t e__ = (call);                                                              \
    if (e__ != cudaSuccess) {                                                              \
      ::sblas::set_error("%s failed at %s:%d : %s", #call, __FILE__, __LINE__,             \
                         cudaGetErrorString(e__));                                         \
      return 1;                                                                            \
    }                                                                                      \
  } while (0)

#define SBLAS_CHECK(cond, ...)                                                             \
  do {                                                                                     \
    if (!(cond)) {                                                                         \
      ::sblas::set_error(__VA_ARGS__);                                                     \
      return 2;                                                                            \
    }                                                                                      \
  } while (0)

// every kernel launch of the library is counted (bench.py reports it as gpu_launches)
extern unsigned long long g_launches;
#define SBLAS_LAUNCH_CHECK()                                                               \
  do {                                                                                     \
    ++::sblas::g_launches;                                                                 \
    cudaError_t e__ = cudaGetLastError();                                                  \
    if (e__ != cudaSuccess) {                                                              \
      ::sblas::set_error("kernel launch failed at %s:%d : %s", __FILE__, __LINE__,         \
                         cudaGetErrorString(e__));                                         \
      return 3;                                                                            \
    }                                                                                      \
  } while (0)

constexpr int kNumSM = 148;  // B200: 2 dies x 74 SMs

// ---- output replicas (fused stitch: the epilogue stores every result to the local y/C
//      and to the same offset of each peer GPU's replica over NVLink) -------------------
template <typename T>
struct OutSpec {
  int n;                        // number of extra replicas
  T* ptr[SBLAS_MAX_GPU - 1];
};

template <typename T>
inline OutSpec<T> make_outspec(const sblasOpts* o) {
  OutSpec<T> s;
  s.n = 0;
  for (int i = 0; i < SBLAS_MAX_GPU - 1; ++i) s.ptr[i] = nullptr;
  if (o) {
    s.n = o->n_replica;
    for (int i = 0; i < o->n_replica && i < SBLAS_MAX_GPU - 1; ++i) s.ptr[i] = (T*)o->replica[i];
  }
  return s;
}

// ---- exactly rounded arithmetic ---------------------------------------------------------
// The reference's CPU loops (spmv.h:26-30, spmm.h:60-66) run as separately rounded
// multiply and add on x86-64 (no FMA in its build flags, Makefile:23).  The sequential
// kernel paths use these so that their results are bit-identical to it; nvcc would
// otherwise contract a*b+c into an FMA.
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// beta*y + alpha*s exactly as `beta * (pC->val[i]) + alpha * sum` (spmv.h:31, spmm.h:66).
template <typename T>
__device__ __forceinline__ T axpby_rn(T alpha, T s, T beta, T y) {
  return add_rn(mul_rn(beta, y), mul_rn(alpha, s));
}

// ---- loads with cache hints ------------------------------------------------------------
// Streaming (touched once) operands: col_idx / val of A.  Bypass L1 so the gathered
// x / B lines keep it.
__device__ __forceinline__ int4 ldg_stream_i4(const int4* p) {
  int4 r;
  asm("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ double2 ldg_stream_d2(const double2* p) {
  double2 r;
  asm("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int ldg_stream(const int* p) {
  int r;
  asm("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ double ldg_stream(const double* p) {
  double r;
  asm("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ float ldg_stream(const float* p) {
  float r;
  asm("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

// L2 eviction policies (createpolicy) and loads that carry one: streamed operands are marked evict_first, gathered
// ones evict_last, so a slice of x that is meant to stay L2-resident is not pushed out by the matrix streaming by
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ int ldg_stream_hint(const int* p, uint64_t pol) {
  int r;
  asm("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ double ldg_stream_hint(const double* p, uint64_t pol) {
  double r;
  asm("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ float ldg_stream_hint(const float* p, uint64_t pol) {
  float r;
  asm("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ double ldg_hint(const double* p, uint64_t pol) {
  double r;
  asm("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ float ldg_hint(const float* p, uint64_t pol) {
  float r;
  asm("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
  return r;
}

// 16-byte vector type of T
template <typename T> struct Vec16;
template <> struct Vec16<double> { using type = double2; static constexpr int N = 2; };
template <> struct Vec16<float> { using type = float4; static constexpr int N = 4; };

__device__ __forceinline__ double2 ldg_v(const double2* p) { return __ldg(p); }
__device__ __forceinline__ float4 ldg_v(const float4* p) { return __ldg(p); }
__device__ __forceinline__ double2 ldg_v_nol1(const double2* p) { return ldg_stream_d2(p); }
__device__ __forceinline__ float4 ldg_v_nol1(const float4* p) { return ldg_stream_f4(p); }

template <typename T>
__device__ __forceinline__ T warp_max(T v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) {
    T o = __shfl_xor_sync(0xffffffffu, v, m);
    v = o > v ? o : v;
  }
  return v;
}

// ---- asynchronous global -> shared staging (sm_90+/sm_100a: TMA 1-D bulk copies, LDGSTS,
//      mbarrier completion).  Used by the pipelined SpMV tile kernel: one thread issues
//      cp.async.bulk (SASS UBLKCP) for the contiguous col_idx / val ranges of a tile, every
//      thread adds small cp.async (LDGSTS) for unaligned heads/tails, row pointers and y,
//      and a single mbarrier phase tracks all of it. ---------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// global -> shared bulk copy; dst, src and bytes are multiples of 16; completion = complete_tx on bar
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// 16 bytes through L1 (.ca): gathered dense rows are re-read by neighbouring block rows
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void prefetch_l1(const void* p) {
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
// 16 bytes past L1 (.cg), destination given as a 32-bit shared address
__device__ __forceinline__ void cp_async16_cg(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void cp_async_elem(int32_t* d, const int32_t* s) { cp_async4(d, s); }
__device__ __forceinline__ void cp_async_elem(float* d, const float* s) { cp_async4(d, s); }
__device__ __forceinline__ void cp_async_elem(double* d, const double* s) { cp_async8(d, s); }
// this thread's earlier cp.async arrive on bar when they land (the count is pre-charged: noinc)
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// shared -> global bulk copy (TMA 1-D store, SASS UBLKCP): dst may be a peer GPU's memory mapped over NVLink;
// dst, src and bytes are multiples of 16; completion is tracked by the issuing thread's bulk async-groups
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed groups of this thread have READ their shared-memory source (it may be overwritten)
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... and have completed (their writes are performed)
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy accesses to shared memory before, async-proxy (bulk copy) writes after
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace sblas
