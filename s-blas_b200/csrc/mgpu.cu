// mgpu.cu — the reference's multi-GPU ops, one process driving GPUs 0..n_gpu-1.
//
// Replaces the bodies of sblas_spmv_csr_v1 (spmv.h:36-139), sblas_spmm_csr_v1 (spmm.h:84-161)
// and sblas_spmm_csr_v2 (spmm.h:164-284).  The reference, per call, creates NCCL
// communicators and cuSPARSE handles, uploads a zero-filled copy of the whole output to every
// GPU, all-reduces that whole buffer and runs an extra axpby pass.  Here:
//   * nothing is created per call except a few events (streams = each device's default stream,
//     so the caller's cudaMemcpy/sync2cpu ordering is the same as with the reference);
//   * owner-computes: GPU i computes the rows of its nnz segment once, with the alpha/beta
//     epilogue fused, and its kernel stores every finished row straight into ALL replicas
//     (local + peers over NVLink/NVSwitch) — the stitch is part of the compute kernel, there
//     is no separate collective and no second pass over the output;
//   * the <= n_gpu-1 rows that the reference's nnz split shares between neighbouring GPUs
//     (matrix.h:443-445) travel as k partial sums each ("carries") and are finished by a
//     tiny kernel on every GPU, together with rows that no segment covers (empty rows
//     outside every [start,stop] range, which the reference scales by beta via its axpby).
#include <vector>

#include "plan.cuh"

namespace sblas {

// NVTX range of the current phase of an op; closed on every exit path
struct PhaseRange {
  explicit PhaseRange(const char* name) { sblas_range_push(name); }
  void next(const char* name) { sblas_range_pop(); sblas_range_push(name); }
  ~PhaseRange() { sblas_range_pop(); }
};

struct FinishList {
  int n_shared;
  int32_t row[SBLAS_MAX_GPU], first[SBLAS_MAX_GPU], last[SBLAS_MAX_GPU];
  int n_gap;
  int32_t gap_lo[SBLAS_MAX_GPU + 1], gap_hi[SBLAS_MAX_GPU + 1];  // uncovered rows [lo, hi)
};

// out[row, n] = beta*out[row, n] + alpha * (carry[first][tail][n] + sum_{g>first} carry[g][head][n])
// and out[r, n] = beta*out[r, n] + alpha*0 for rows no segment covers.
// tab[g] = the carry table ([gpu][head|tail][kpad]) in which GPU g's carries are found: with peer access GPU g's OWN
// table, read over NVLink by the finish kernel of every other GPU (a few hundred bytes; this removes the n*(n-1)
// cudaMemcpyPeerAsync calls per op that cost 0.5 ms of host time at 8 GPUs); else the local copy of it
template <typename T>
struct CarryTabs { const T* tab[SBLAS_MAX_GPU]; };

template <typename T>
__global__ void __launch_bounds__(256)
finish_rows_kernel(FinishList fl, CarryTabs<T> ct, int32_t kpad, int32_t k,
                   int32_t layout, T* out, int64_t ld, T alpha, T beta) {
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = tid; e < (int64_t)fl.n_shared * k; e += nth) {
    const int s = (int)(e / k), n = (int)(e % k);
    T sum = ct.tab[fl.first[s]][((int64_t)fl.first[s] * 2 + 1) * kpad + n];
    for (int g = fl.first[s] + 1; g <= fl.last[s]; ++g) sum += ct.tab[g][((int64_t)g * 2) * kpad + n];
    const int64_t off = layout == SBLAS_ROW_MAJOR ? (int64_t)fl.row[s] * ld + n : (int64_t)n * ld + fl.row[s];
    out[off] = axpby_rn(alpha, sum, beta, out[off]);
  }
  for (int gidx = 0; gidx < fl.n_gap; ++gidx) {
    const int64_t lo = fl.gap_lo[gidx], cnt = (int64_t)fl.gap_hi[gidx] - lo;
    for (int64_t e = tid; e < cnt * k; e += nth) {
      int64_t r, n;
      if (layout == SBLAS_ROW_MAJOR) { r = lo + e / k; n = e % k; } else { n = e / cnt; r = lo + e % cnt; }
      const int64_t off = layout == SBLAS_ROW_MAJOR ? r * ld + n : n * ld + r;
      out[off] = axpby_rn(alpha, T(0), beta, out[off]);
    }
  }
}

template <typename T>
static int launch_finish(cudaStream_t st, const FinishList& fl, const CarryTabs<T>& ct, int32_t kpad,
                         int32_t k, int32_t layout, T* out, int64_t ld, T alpha, T beta) {
  int64_t work = (int64_t)fl.n_shared * k;
  for (int g = 0; g < fl.n_gap; ++g) work = std::max<int64_t>(work, ((int64_t)fl.gap_hi[g] - fl.gap_lo[g]) * k);
  if (work <= 0) return 0;
  const int blocks = (int)std::min<int64_t>(ceil_div(work, 256), kNumSM * 8);
  finish_rows_kernel<T><<<blocks, 256, 0, st>>>(fl, ct, kpad, k, layout, out, ld, alpha, beta);
  SBLAS_LAUNCH_CHECK();
  return 0;
}

// ---- device-side barrier between the ranks of one node (one process per GPU) -------------------
// Completes a peer-store stitched step without any host synchronisation: thread t of rank r
// release-stores `epoch` into slot r of rank t's flag array (over NVLink), then spins (acquire,
// system scope) until slot t of its own array reaches `epoch`.  The stores a rank's SpMV/SpMM
// kernel made into the peers' replicas precede its flag store in stream order, so when the wait
// returns every peer's rows have landed in the local replica.  Each rank owns one GPU, so the
// spinning kernels of different ranks always run concurrently; the spin is bounded (20 s of
// %globaltimer by default, SBLAS_BARRIER_TIMEOUT_MS) and reports a timeout through the sticky
// *status word — which RowSplitOp.check() turns into an exception — instead of hanging the device.
__global__ void peer_barrier_kernel(int rank, int world, int32_t* local_flags, OutSpec<int32_t> peers,
                                    int32_t epoch, int32_t* status, unsigned long long timeout_ns, int trap) {
  const int t = threadIdx.x;
  if (t >= world) return;
  if (t != rank) {
    int32_t* dst = peers.ptr[t] + rank;
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
    unsigned long long t0, now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    int32_t v;
    for (;;) {
      asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(local_flags + t) : "memory");
      if (v - epoch >= 0) break;
      __nanosleep(64);
      if (timeout_ns) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > timeout_ns) break;
      }
    }
    if (v - epoch < 0) {   // the replica is INCOMPLETE: say so (sticky), and fail the stream when asked to
      if (status) atomicExch(status, 1 + t);
      if (trap) __trap();
    }
  }
}

// wait budget of the barrier: SBLAS_BARRIER_TIMEOUT_MS (default 20 000; 0 = wait for ever, e.g. under a
// profiler that serialises the ranks), SBLAS_BARRIER_TRAP=1 turns a timeout into a device trap (the stream
// then reports an error on its next synchronisation instead of only setting the status word)
static void barrier_policy(unsigned long long* timeout_ns, int* trap) {
  static const long long ms = [] { const char* e = getenv("SBLAS_BARRIER_TIMEOUT_MS"); return e ? atoll(e) : 20000ll; }();
  static const int tr = [] { const char* e = getenv("SBLAS_BARRIER_TRAP"); return e ? atoi(e) : 0; }();
  *timeout_ns = ms > 0 ? (unsigned long long)ms * 1000000ull : 0ull;
  *trap = tr;
}

// ---- logical GPU -> physical device ------------------------------------------------------------
// SBLAS_DEVICE_MAP="0,0,1,1" runs a 4-way partition on two physical devices (several logical
// GPUs may share one device: their kernels then simply run back to back on its default
// stream).  Default: identity, i.e. the reference's cudaSetDevice(i_gpu) (spmv.h:55).
static int g_devmap[SBLAS_MAX_GPU];
static bool g_devmap_init = false;
static int device_of(int i) {
  if (!g_devmap_init) {
    for (int j = 0; j < SBLAS_MAX_GPU; ++j) g_devmap[j] = j;
    if (const char* e = getenv("SBLAS_DEVICE_MAP")) {
      int j = 0;
      while (*e && j < SBLAS_MAX_GPU) {
        g_devmap[j++] = (int)strtol(e, (char**)&e, 10);
        while (*e == ',' || *e == ' ') ++e;
      }
    }
    g_devmap_init = true;
  }
  return (i >= 0 && i < SBLAS_MAX_GPU) ? g_devmap[i] : i;
}

// ---- cached context -----------------------------------------------------------------------------
struct MgpuCtx {
  int n = 0;
  bool peer_ok = false;
  void* carry[SBLAS_MAX_GPU] = {nullptr};  // per GPU: [n][2][kpad] elements of 8 bytes max
  size_t carry_bytes = 0;
  cudaEvent_t ev[SBLAS_MAX_GPU] = {nullptr};
};
static MgpuCtx g_ctx;

static int ctx_get(unsigned n_gpu, size_t carry_bytes, MgpuCtx** out) {
  SBLAS_CHECK(n_gpu >= 1 && n_gpu <= SBLAS_MAX_GPU, "n_gpu must be in 1..%d", SBLAS_MAX_GPU);
  int ndev = 0;
  SBLAS_CUDA(cudaGetDeviceCount(&ndev));
  for (unsigned i = 0; i < n_gpu; ++i)
    SBLAS_CHECK(device_of(i) < ndev, "logical GPU %u maps to device %d but only %d CUDA devices are visible", i, device_of(i), ndev);
  int cur = 0;
  SBLAS_CUDA(cudaGetDevice(&cur));
  if (g_ctx.n < (int)n_gpu) {
    bool ok = true;
    for (unsigned i = 0; i < n_gpu; ++i) {
      SBLAS_CUDA(cudaSetDevice(device_of(i)));
      for (unsigned j = 0; j < n_gpu; ++j) {
        if (device_of(i) == device_of(j)) continue;
        int can = 0;
        SBLAS_CUDA(cudaDeviceCanAccessPeer(&can, device_of(i), device_of(j)));
        if (!can) { ok = false; continue; }
        cudaError_t e = cudaDeviceEnablePeerAccess(device_of(j), 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) { cudaGetLastError(); ok = false; }
      }
      if (!g_ctx.ev[i]) SBLAS_CUDA(cudaEventCreateWithFlags(&g_ctx.ev[i], cudaEventDisableTiming));
    }
    g_ctx.peer_ok = ok;
    g_ctx.n = n_gpu;
  }
  if (g_ctx.carry_bytes < carry_bytes) {
    for (int i = 0; i < SBLAS_MAX_GPU; ++i)
      if (g_ctx.carry[i]) { cudaSetDevice(device_of(i)); cudaFree(g_ctx.carry[i]); g_ctx.carry[i] = nullptr; }
    g_ctx.carry_bytes = 0;
  }
  for (unsigned i = 0; i < n_gpu; ++i)
    if (!g_ctx.carry[i]) {
      SBLAS_CUDA(cudaSetDevice(device_of(i)));
      SBLAS_CUDA(cudaMalloc(&g_ctx.carry[i], std::max<size_t>(carry_bytes, 4096)));
    }
  g_ctx.carry_bytes = std::max<size_t>(std::max<size_t>(carry_bytes, 4096), g_ctx.carry_bytes);
  SBLAS_CUDA(cudaSetDevice(cur));
  *out = &g_ctx;
  return 0;
}

// Which rows are shared / uncovered, from the reference's partition tables alone.
static void build_finish(unsigned n, int32_t height, const int32_t* nnz_gpu, const int32_t* s,
                         const int32_t* e, bool* active, bool* head, bool* tail, FinishList* fl) {
  fl->n_shared = 0;
  fl->n_gap = 0;
  for (unsigned i = 0; i < n; ++i) {
    active[i] = nnz_gpu[i] > 0 && s[i] >= 0 && e[i] >= s[i];
    head[i] = tail[i] = false;
  }
  int prev = -1;       // previous active GPU
  int32_t covered = 0; // rows [0, covered) are covered so far
  for (unsigned i = 0; i < n; ++i) {
    if (!active[i]) continue;
    if (prev >= 0 && e[prev] == s[i]) {  // row s[i] continues from GPU prev
      head[i] = true;
      tail[prev] = true;
      if (fl->n_shared > 0 && fl->row[fl->n_shared - 1] == s[i]) fl->last[fl->n_shared - 1] = i;
      else { fl->row[fl->n_shared] = s[i]; fl->first[fl->n_shared] = prev; fl->last[fl->n_shared] = i; fl->n_shared++; }
    }
    if (s[i] > covered) { fl->gap_lo[fl->n_gap] = covered; fl->gap_hi[fl->n_gap] = s[i]; fl->n_gap++; }
    covered = e[i] + 1;
    prev = i;
  }
  if (covered < height) { fl->gap_lo[fl->n_gap] = covered; fl->gap_hi[fl->n_gap] = height; fl->n_gap++; }
}

// SpMV (k = 1, layout ignored) and SpMM method 2 share this driver.
template <typename T>
static int mgpu_rowsplit(unsigned n_gpu, int32_t height, int32_t width,
                         const int32_t* const* rp, const int32_t* const* ci, const T* const* val,
                         const int32_t* nnz_gpu, const int32_t* srow, const int32_t* erow,
                         bool is_spmm, int32_t k, int32_t layout, const T* const* B, int64_t ldb,
                         T* const* C, int64_t ldc, T alpha, T beta, const sblasPlan_t* plans) {
  SBLAS_CHECK(rp && ci && val && nnz_gpu && srow && erow && B && C, "mgpu: null argument");
  if (!is_spmm) { k = 1; layout = SBLAS_COL_MAJOR; ldb = width; ldc = height; }
  const int32_t kpad = (k + 3) & ~3;
  MgpuCtx* ctx = nullptr;
  if (ctx_get(n_gpu, (size_t)n_gpu * 2 * kpad * sizeof(double), &ctx)) return 1;
  int cur = 0;
  SBLAS_CUDA(cudaGetDevice(&cur));
  bool active[SBLAS_MAX_GPU], head[SBLAS_MAX_GPU], tail[SBLAS_MAX_GPU];
  FinishList fl;
  build_finish(n_gpu, height, nnz_gpu, srow, erow, active, head, tail, &fl);
  const bool peer = ctx->peer_ok && n_gpu > 1;
  int rc = 0;

  // 1. every GPU: its segment, fused epilogue, stores into all replicas
  PhaseRange phase("sblas mgpu: segment kernels + fused peer stitch");
  for (unsigned i = 0; i < n_gpu && rc == 0; ++i) {
    if (!active[i]) continue;
    SBLAS_CUDA(cudaSetDevice(device_of(i)));
    const int32_t rows = erow[i] - srow[i] + 1;
    sblasOpts o;
    memset(&o, 0, sizeof o);
    o.flags = (head[i] ? SBLAS_FLAG_HEAD_PARTIAL : 0) | (tail[i] ? SBLAS_FLAG_TAIL_PARTIAL : 0);
    T* carry_i = (T*)ctx->carry[i] + (size_t)i * 2 * kpad;
    o.carry = carry_i;
    const int64_t row_off = (is_spmm && layout == SBLAS_ROW_MAJOR) ? (int64_t)srow[i] * ldc : srow[i];
    if (peer) {
      for (unsigned d = 0; d < n_gpu; ++d)
        if (d != i) o.replica[o.n_replica++] = C[d] + row_off;
    }
    sblasPlan_t plan = plans ? plans[i] : nullptr;
    if (is_spmm)
      rc = sizeof(T) == 8
               ? sblas_spmm_csr_f64(nullptr, rows, width, nnz_gpu[i], rp[i], ci[i], (const double*)val[i], k, layout, (const double*)B[i], ldb, (double*)(C[i] + row_off), ldc, alpha, beta, plan, &o)
               : sblas_spmm_csr_f32(nullptr, rows, width, nnz_gpu[i], rp[i], ci[i], (const float*)val[i], k, layout, (const float*)B[i], ldb, (float*)(C[i] + row_off), ldc, alpha, beta, plan, &o);
    else
      rc = sizeof(T) == 8
               ? sblas_spmv_csr_f64(nullptr, rows, width, nnz_gpu[i], rp[i], ci[i], (const double*)val[i], (const double*)B[i], (double*)(C[i] + row_off), alpha, beta, plan, &o)
               : sblas_spmv_csr_f32(nullptr, rows, width, nnz_gpu[i], rp[i], ci[i], (const float*)val[i], (const float*)B[i], (float*)(C[i] + row_off), alpha, beta, plan, &o);
    if (rc) break;
    // carries of GPU i to every other GPU's carry table (tiny) — only without peer access: with it the finish
    // kernels read them in place
    if (!peer && (head[i] || tail[i]))
      for (unsigned d = 0; d < n_gpu; ++d)
        if (d != i)
          SBLAS_CUDA(cudaMemcpyPeerAsync((T*)ctx->carry[d] + (size_t)i * 2 * kpad, device_of(d), carry_i, device_of(i),
                                         sizeof(T) * 2 * kpad, nullptr));
    // without peer stores: ship the finished panel to the other replicas
    if (!peer && n_gpu > 1) {
      for (unsigned d = 0; d < n_gpu; ++d) {
        if (d == i) continue;
        if (!is_spmm || layout == SBLAS_ROW_MAJOR) {
          const size_t bytes = sizeof(T) * (is_spmm ? (size_t)rows * ldc : (size_t)rows);
          SBLAS_CUDA(cudaMemcpyPeerAsync(C[d] + row_off, device_of(d), C[i] + row_off, device_of(i), bytes, nullptr));
        } else {
          for (int32_t n = 0; n < k; ++n)
            SBLAS_CUDA(cudaMemcpyPeerAsync(C[d] + row_off + (int64_t)n * ldc, device_of(d), C[i] + row_off + (int64_t)n * ldc, device_of(i), sizeof(T) * rows, nullptr));
        }
      }
    }
    SBLAS_CUDA(cudaEventRecord(ctx->ev[i], nullptr));
  }
  phase.next("sblas mgpu: carries + finish rows + device sync");
  // 2. every GPU: wait for all segments (and carries / panel copies), then finish the shared
  //    and the uncovered rows of its own replica.  Step 1 never writes those rows with new
  //    values (a panel copy can only carry their old value), so running this last is enough.
  for (unsigned d = 0; d < n_gpu && rc == 0; ++d) {
    SBLAS_CUDA(cudaSetDevice(device_of(d)));
    for (unsigned i = 0; i < n_gpu; ++i)
      if (i != d && active[i]) SBLAS_CUDA(cudaStreamWaitEvent(nullptr, ctx->ev[i], 0));
    CarryTabs<T> ct;
    for (unsigned g = 0; g < SBLAS_MAX_GPU; ++g) ct.tab[g] = (const T*)ctx->carry[(peer && g < n_gpu) ? g : d];
    rc = launch_finish<T>(nullptr, fl, ct, kpad, k, layout, C[d], ldc, alpha, beta);
  }
  for (unsigned d = 0; d < n_gpu; ++d) {
    cudaSetDevice(device_of(d));
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess && rc == 0) { set_error("mgpu: device %u: %s", d, cudaGetErrorString(e)); rc = 1; }
  }
  cudaSetDevice(cur);
  return rc;
}

template <typename T>
static int mgpu_colsplit(unsigned n_gpu, int32_t height, int32_t width, int32_t nnz,
                         const int32_t* const* rp, const int32_t* const* ci, const T* const* val,
                         const int32_t* dim_gpu, int32_t layout, const T* const* B,
                         const int64_t* ldb, T* const* C, const int64_t* ldc, T alpha, T beta,
                         const sblasPlan_t* plans) {
  SBLAS_CHECK(rp && ci && val && dim_gpu && B && C && ldb && ldc, "mgpu: null argument");
  SBLAS_CHECK(n_gpu >= 1 && n_gpu <= SBLAS_MAX_GPU, "n_gpu must be in 1..%d", SBLAS_MAX_GPU);
  int cur = 0;
  SBLAS_CUDA(cudaGetDevice(&cur));
  int rc = 0;
  for (unsigned i = 0; i < n_gpu && rc == 0; ++i) {
    if (dim_gpu[i] <= 0) continue;  // trailing GPUs may get nothing
    SBLAS_CUDA(cudaSetDevice(device_of(i)));
    sblasPlan_t plan = plans ? plans[i] : nullptr;
    rc = sizeof(T) == 8
             ? sblas_spmm_csr_f64(nullptr, height, width, nnz, rp[i], ci[i], (const double*)val[i], dim_gpu[i], layout, (const double*)B[i], ldb[i], (double*)C[i], ldc[i], alpha, beta, plan, nullptr)
             : sblas_spmm_csr_f32(nullptr, height, width, nnz, rp[i], ci[i], (const float*)val[i], dim_gpu[i], layout, (const float*)B[i], ldb[i], (float*)C[i], ldc[i], alpha, beta, plan, nullptr);
  }
  for (unsigned d = 0; d < n_gpu; ++d) {
    cudaSetDevice(device_of(d));
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess && rc == 0) { set_error("mgpu: device %u: %s", d, cudaGetErrorString(e)); rc = 1; }
  }
  cudaSetDevice(cur);
  return rc;
}

}  // namespace sblas

using namespace sblas;
extern "C" {

int sblas_device_of(int logical_gpu) { return device_of(logical_gpu); }

// ---- CUDA-IPC shared replicas (one process per GPU: torchrun / torch.distributed) ------------
int sblas_ipc_alloc(size_t bytes, void** ptr, unsigned char handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  SBLAS_CHECK(ptr && handle && bytes > 0, "ipc_alloc: bad arguments");
  SBLAS_CUDA(cudaMalloc(ptr, bytes));
  cudaIpcMemHandle_t h;
  SBLAS_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle, &h, 64);
  return 0;
}
int sblas_ipc_open(const unsigned char handle[64], void** ptr) {
  SBLAS_CHECK(ptr && handle, "ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  SBLAS_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int sblas_ipc_close(void* ptr) {
  if (ptr) SBLAS_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
int sblas_ipc_free(void* ptr) {
  if (ptr) SBLAS_CUDA(cudaFree(ptr));
  return 0;
}

int sblas_peer_barrier(void* stream, int rank, int world, int32_t* d_local_flags,
                       void* const* d_peer_flags, int32_t epoch, int32_t* d_status) {
  SBLAS_CHECK(world >= 1 && world <= SBLAS_MAX_GPU && rank >= 0 && rank < world, "peer_barrier: bad rank/world");
  SBLAS_CHECK(d_local_flags && d_peer_flags, "peer_barrier: null flags");
  if (world == 1) return 0;
  OutSpec<int32_t> peers;
  peers.n = world;
  for (int i = 0; i < SBLAS_MAX_GPU - 1; ++i) peers.ptr[i] = nullptr;
  SBLAS_CHECK(world <= SBLAS_MAX_GPU - 1, "peer_barrier: at most %d ranks", SBLAS_MAX_GPU - 1);
  for (int i = 0; i < world; ++i) peers.ptr[i] = (int32_t*)d_peer_flags[i];
  unsigned long long timeout_ns = 0;
  int trap = 0;
  barrier_policy(&timeout_ns, &trap);
  peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(rank, world, d_local_flags, peers, epoch, d_status, timeout_ns, trap);
  SBLAS_LAUNCH_CHECK();
  return 0;
}

#define SBLAS_MGPU_SPMV(SFX, T)                                                                    \
  int sblas_mgpu_spmv_csr_v1_##SFX(unsigned n_gpu, int32_t height, int32_t width,                  \
                                   const int32_t* const* rp, const int32_t* const* ci,             \
                                   const T* const* val, const int32_t* nnz_gpu,                    \
                                   const int32_t* srow, const int32_t* erow, const T* const* x,    \
                                   T* const* y, T alpha, T beta, const sblasPlan_t* plans) {       \
    return mgpu_rowsplit<T>(n_gpu, height, width, rp, ci, val, nnz_gpu, srow, erow, false, 1,      \
                            SBLAS_COL_MAJOR, x, width, y, height, alpha, beta, plans);             \
  }
SBLAS_MGPU_SPMV(f64, double)
SBLAS_MGPU_SPMV(f32, float)

#define SBLAS_MGPU_SPMM(SFX, T)                                                                    \
  int sblas_mgpu_spmm_csr_v1_##SFX(unsigned n_gpu, int32_t height, int32_t width, int32_t nnz,     \
                                   const int32_t* const* rp, const int32_t* const* ci,             \
                                   const T* const* val, const int32_t* dim_gpu, int32_t layout,    \
                                   const T* const* B, const int64_t* ldb, T* const* C,             \
                                   const int64_t* ldc, T alpha, T beta,                            \
                                   const sblasPlan_t* plans) {                                     \
    return mgpu_colsplit<T>(n_gpu, height, width, nnz, rp, ci, val, dim_gpu, layout, B, ldb, C,    \
                            ldc, alpha, beta, plans);                                              \
  }                                                                                                \
  int sblas_mgpu_spmm_csr_v2_##SFX(unsigned n_gpu, int32_t height, int32_t width,                  \
                                   const int32_t* const* rp, const int32_t* const* ci,             \
                                   const T* const* val, const int32_t* nnz_gpu,                    \
                                   const int32_t* srow, const int32_t* erow, int32_t k,            \
                                   int32_t layout, const T* const* B, int64_t ldb, T* const* C,    \
                                   int64_t ldc, T alpha, T beta, const sblasPlan_t* plans) {       \
    return mgpu_rowsplit<T>(n_gpu, height, width, rp, ci, val, nnz_gpu, srow, erow, true, k,       \
                            layout, B, ldb, C, ldc, alpha, beta, plans);                           \
  }
SBLAS_MGPU_SPMM(f64, double)
SBLAS_MGPU_SPMM(f32, float)

#define SBLAS_FIXUP(SFX, T)                                                                        \
  int sblas_carry_fixup_##SFX(void* stream, int32_t n_shared, const int32_t* rows,                 \
                              const int32_t* first, const int32_t* last, const T* d_carry,         \
                              int32_t k, int32_t layout, T* d_out, int64_t ld, T alpha, T beta) {  \
    SBLAS_CHECK(n_shared >= 0 && n_shared <= SBLAS_MAX_GPU, "carry_fixup: too many shared rows");  \
    FinishList fl;                                                                                 \
    fl.n_gap = 0;                                                                                  \
    fl.n_shared = n_shared;                                                                        \
    for (int i = 0; i < n_shared; ++i) { fl.row[i] = rows[i]; fl.first[i] = first[i]; fl.last[i] = last[i]; } \
    CarryTabs<T> ct;                                                                               \
    for (int g = 0; g < SBLAS_MAX_GPU; ++g) ct.tab[g] = d_carry;                                   \
    return launch_finish<T>((cudaStream_t)stream, fl, ct, (k + 3) & ~3, k, layout, d_out, ld,      \
                            alpha, beta);                                                          \
  }
SBLAS_FIXUP(f64, double)
SBLAS_FIXUP(f32, float)

}  // extern "C"
