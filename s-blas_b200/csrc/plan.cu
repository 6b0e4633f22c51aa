// plan.cu — CSR row statistics + merge-path tile tables (see plan.cuh).
#include <stdarg.h>

#include <mutex>

#include "plan.cuh"

namespace sblas {

unsigned long long g_launches = 0;
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

// stats[0] = max row nnz, stats[1] = empty rows
__global__ void row_stats_kernel(const int32_t* __restrict__ rowptr, int32_t rows, int32_t* stats) {
  int32_t mx = 0, empty = 0;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows;
       r += (int64_t)gridDim.x * blockDim.x) {
    int32_t len = rowptr[r + 1] - rowptr[r];
    mx = max(mx, len);
    empty += (len == 0);
  }
  mx = warp_max(mx);
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) empty += __shfl_xor_sync(0xffffffffu, empty, m);
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&stats[0], mx);
    atomicAdd(&stats[1], empty);
  }
}

// tile_row[b] = smallest r in [0, rows] with r*rw + rowptr[r] >= b*T  (b = 0..n_tiles)
__global__ void tile_search_kernel(const int32_t* __restrict__ rowptr, int32_t rows, int32_t rw,
                                   int32_t T, int32_t n_tiles, int32_t* __restrict__ tile_row,
                                   int32_t* __restrict__ tile_nnz) {
  int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > n_tiles) return;
  if (b == n_tiles) {
    tile_row[b] = rows;
    tile_nnz[b] = rowptr[rows];
    return;
  }
  int64_t target = (int64_t)b * T;
  int32_t lo = 0, hi = rows;  // answer in [lo, hi]
  while (lo < hi) {
    int32_t mid = lo + ((hi - lo) >> 1);
    int64_t f = (int64_t)mid * rw + rowptr[mid];
    if (f >= target) hi = mid; else lo = mid + 1;
  }
  tile_row[b] = lo;
  tile_nnz[b] = rowptr[lo];
}

// stats[0] = max nnz of a tile, stats[1] = max rows of a tile
__global__ void tile_stats_kernel(const int32_t* __restrict__ rowptr,
                                  const int32_t* __restrict__ tile_row, int32_t n_tiles,
                                  int32_t* stats) {
  int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  int32_t nz = 0, nr = 0;
  if (b < n_tiles) {
    int32_t r0 = tile_row[b], r1 = tile_row[b + 1];
    nr = r1 - r0;
    nz = rowptr[r1] - rowptr[r0];
  }
  nz = warp_max(nz);
  nr = warp_max(nr);
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&stats[0], nz);
    atomicMax(&stats[1], nr);
  }
}

namespace {
struct KernInfo { const void* kern; int device, block; size_t smem; int per_sm, n_sm; };
struct KernAttr { const void* kern; int device; size_t max_smem; };
std::mutex g_kern_mu;
std::vector<KernInfo> g_kern_info;
std::vector<KernAttr> g_kern_attr;
}  // namespace

int kernel_residency(const void* kern, int block, size_t smem, int* per_sm, int* n_sm) {
  int device = 0;
  SBLAS_CUDA(cudaGetDevice(&device));
  std::lock_guard<std::mutex> lock(g_kern_mu);
  for (const KernInfo& k : g_kern_info)
    if (k.kern == kern && k.device == device && k.block == block && k.smem == smem) {
      *per_sm = k.per_sm;
      *n_sm = k.n_sm;
      return 0;
    }
  if (smem > 48 * 1024) {  // opt-in limit: per function and device, only ever raised
    SBLAS_CHECK(smem <= 227 * 1024, "kernel needs %zu bytes of shared memory", smem);
    KernAttr* a = nullptr;
    for (KernAttr& x : g_kern_attr)
      if (x.kern == kern && x.device == device) a = &x;
    if (!a || a->max_smem < smem) {
      SBLAS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (a) a->max_smem = smem; else g_kern_attr.push_back({kern, device, smem});
    }
  }
  KernInfo k{kern, device, block, smem, 0, kNumSM};
  SBLAS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&k.per_sm, kern, block, smem));
  SBLAS_CHECK(k.per_sm >= 1, "kernel does not fit an SM with %zu bytes of shared memory", smem);
  SBLAS_CUDA(cudaDeviceGetAttribute(&k.n_sm, cudaDevAttrMultiProcessorCount, device));
  g_kern_info.push_back(k);
  *per_sm = k.per_sm;
  *n_sm = k.n_sm;
  return 0;
}

int plan_get_table(sblasPlan_st* plan, cudaStream_t stream, int32_t rw, int32_t T,
                   const TileTable** out) {
  for (const TileTable& t : plan->tables)
    if (t.tile == T && t.row_weight == rw) {
      *out = &t;
      return 0;
    }
  SBLAS_CHECK(T > 0 && rw > 0, "plan: bad tile parameters");
  TileTable t;
  t.tile = T;
  t.row_weight = rw;
  int64_t items = (int64_t)plan->rows * rw + plan->nnz;
  int64_t nt = ceil_div(items > 0 ? items : 1, T);
  SBLAS_CHECK(nt < (1ll << 31) - 2, "plan: too many tiles");
  t.n_tiles = (int32_t)nt;
  int32_t* d_stats = nullptr;
  int32_t h_stats[2] = {0, 0};
  const int blocks = (int)ceil_div(t.n_tiles + 1, 256);
  auto build = [&]() -> int {  // any failure below releases the partial allocations
    SBLAS_CUDA(cudaMalloc(&t.d_tile_row, sizeof(int32_t) * (size_t)(t.n_tiles + 1)));
    SBLAS_CUDA(cudaMalloc(&t.d_tile_nnz, sizeof(int32_t) * (size_t)(t.n_tiles + 1)));
    SBLAS_CUDA(cudaMalloc(&d_stats, 2 * sizeof(int32_t)));
    SBLAS_CUDA(cudaMemsetAsync(d_stats, 0, 2 * sizeof(int32_t), stream));
    tile_search_kernel<<<blocks, 256, 0, stream>>>(plan->d_rowptr, plan->rows, rw, T, t.n_tiles,
                                                   t.d_tile_row, t.d_tile_nnz);
    SBLAS_LAUNCH_CHECK();
    tile_stats_kernel<<<blocks, 256, 0, stream>>>(plan->d_rowptr, t.d_tile_row, t.n_tiles, d_stats);
    SBLAS_LAUNCH_CHECK();
    SBLAS_CUDA(cudaMemcpyAsync(h_stats, d_stats, sizeof h_stats, cudaMemcpyDeviceToHost, stream));
    SBLAS_CUDA(cudaStreamSynchronize(stream));
    return 0;
  };
  const int brc = build();
  if (d_stats) cudaFree(d_stats);
  if (brc) {
    if (t.d_tile_row) cudaFree(t.d_tile_row);
    if (t.d_tile_nnz) cudaFree(t.d_tile_nnz);
    return brc;
  }
  t.max_tile_nnz = h_stats[0];
  t.max_tile_rows = h_stats[1];
  plan->tables.reserve(16);  // pointers handed out stay valid for the first 16 tables
  plan->tables.push_back(t);
  *out = &plan->tables.back();
  return 0;
}

int plan_scratch(sblasPlan_st* plan, size_t bytes, void** out) {
  if (plan->split_scratch_bytes < bytes) {
    if (plan->d_split_scratch) SBLAS_CUDA(cudaFree(plan->d_split_scratch));
    plan->d_split_scratch = nullptr;
    plan->split_scratch_bytes = 0;
    SBLAS_CUDA(cudaMalloc(&plan->d_split_scratch, bytes));
    plan->split_scratch_bytes = bytes;
  }
  *out = plan->d_split_scratch;
  return 0;
}

}  // namespace sblas

using namespace sblas;

extern "C" {

int sblas_version(void) { return 100; }
unsigned long long sblas_launch_count(void) { return g_launches; }
const char* sblas_last_error(void) { return get_error(); }
int sblas_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int sblas_plan_create_i32(void* stream_, int32_t rows, int64_t nnz, const int32_t* d_rowptr,
                          sblasPlan_t* out) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SBLAS_CHECK(out != nullptr && rows >= 0 && nnz >= 0, "plan_create: bad arguments");
  SBLAS_CHECK(d_rowptr != nullptr, "plan_create: null rowptr");
  sblasPlan_st* p = new sblasPlan_st();
  SBLAS_CUDA(cudaGetDevice(&p->device));
  p->rows = rows;
  p->nnz = nnz;
  p->d_rowptr = d_rowptr;
  int32_t* d_stats = nullptr;
  SBLAS_CUDA(cudaMalloc(&d_stats, 2 * sizeof(int32_t)));
  SBLAS_CUDA(cudaMemsetAsync(d_stats, 0, 2 * sizeof(int32_t), stream));
  if (rows > 0) {
    int blocks = (int)std::min<int64_t>(ceil_div(rows, 256), kNumSM * 8);
    row_stats_kernel<<<blocks, 256, 0, stream>>>(d_rowptr, rows, d_stats);
    SBLAS_LAUNCH_CHECK();
  }
  int32_t h[2] = {0, 0};
  SBLAS_CUDA(cudaMemcpyAsync(h, d_stats, sizeof h, cudaMemcpyDeviceToHost, stream));
  SBLAS_CUDA(cudaStreamSynchronize(stream));
  SBLAS_CUDA(cudaFree(d_stats));
  p->max_row_nnz = h[0];
  p->empty_rows = h[1];
  *out = p;
  return 0;
}

int sblas_plan_destroy(sblasPlan_t p) {
  if (!p) return 0;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(p->device);
  for (TileTable& t : p->tables) { cudaFree(t.d_tile_row); if (t.d_tile_nnz) cudaFree(t.d_tile_nnz); }
  if (p->d_split_scratch) cudaFree(p->d_split_scratch);
  rowblocks_free(p);
  colpanels_free(p);
  cudaSetDevice(cur);
  delete p;
  return 0;
}

}  // extern "C"
