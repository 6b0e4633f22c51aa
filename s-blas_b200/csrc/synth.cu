// synth.cu — deterministic synthetic inputs for BASELINE.json's configs, generated on the
// device (the reference fetches SuiteSparse matrices over the network,
// matrices/get_matrices.sh:4-5; there is no network here).  Counter-based: every value is a
// pure function of (seed, stream, index), restated in numpy in tests/synth_ref.py.
#include <cub/cub.cuh>

#include "common.cuh"

namespace sblas {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint64_t hash3(uint64_t seed, uint64_t stream, uint64_t idx) {
  return splitmix64(splitmix64(seed ^ (stream * 0xD1342543DE82EF95ull)) + idx);
}
// u in [0,1): 53 bits for double, 24 bits for float (both exact, so numpy reproduces them)
__device__ __forceinline__ double u01_f64(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }
__device__ __forceinline__ float u01_f32(uint64_t h) { return (float)(h >> 40) * (1.0f / 16777216.0f); }

__device__ __forceinline__ void put_val(void* val, int64_t i, bool f64, uint64_t h, double lo, double hi) {
  if (f64) ((double*)val)[i] = __dadd_rn(lo, __dmul_rn(hi - lo, u01_f64(h)));
  else ((float*)val)[i] = __fadd_rn((float)lo, __fmul_rn((float)hi - (float)lo, u01_f32(h)));
}

template <typename T>
__global__ void fill_kernel(T* p, int64_t n, uint64_t seed, uint64_t stream, T lo, T hi) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    put_val(p, i, sizeof(T) == 8, hash3(seed, stream, (uint64_t)i), (double)lo, (double)hi);
}

// ---- 2-D 5-point Laplacian ---------------------------------------------------------------------
// nnz before row i = i + [rows<i with y>0] + [x>0] + [x<nx-1] + [y<ny-1], all closed form.
__device__ __forceinline__ int64_t lap_prefix(int64_t i, int64_t nx, int64_t ny) {
  const int64_t y = i / nx, x = i % nx;
  int64_t p = i;
  p += i > nx ? i - nx : 0;                       // up neighbour (y > 0)
  p += y * (nx - 1) + (x > 0 ? x - 1 : 0);        // left neighbour (x > 0)
  p += y * (nx - 1) + (x < nx - 1 ? x : nx - 1);  // right neighbour (x < nx-1)
  p += i < (ny - 1) * nx ? i : (ny - 1) * nx;     // down neighbour (y < ny-1)
  return p;
}
__global__ void laplace_kernel(int32_t nx, int32_t ny, int32_t* rowptr, int32_t* col, void* val, bool f64) {
  const int64_t n = (int64_t)nx * ny;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i <= n;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = lap_prefix(i, nx, ny);
    rowptr[i] = (int32_t)p;
    if (i == n) continue;
    const int64_t y = i / nx, x = i % nx;
    auto put = [&](int64_t c, double v) {
      col[p] = (int32_t)c;
      if (f64) ((double*)val)[p] = v; else ((float*)val)[p] = (float)v;
      ++p;
    };
    if (y > 0) put(i - nx, -1.0);
    if (x > 0) put(i - 1, -1.0);
    put(i, 4.0);
    if (x < nx - 1) put(i + 1, -1.0);
    if (y < ny - 1) put(i + nx, -1.0);
  }
}

// ---- uniform random, fixed degree -----------------------------------------------------------------
constexpr int kMaxDeg = 64;
__global__ void uniform_kernel(int32_t rows, int32_t cols, int32_t deg, uint64_t seed,
                               int32_t* rowptr, int32_t* col, void* val, bool f64) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r <= rows;
       r += (int64_t)gridDim.x * blockDim.x) {
    rowptr[r] = (int32_t)(r * deg);
    if (r == rows) continue;
    int32_t c[kMaxDeg];
    int n = 0;
    // draw until `deg` distinct columns; keep them sorted by insertion
    for (uint64_t t = 0; n < deg; ++t) {
      const int32_t x = (int32_t)(hash3(seed, 2 * (uint64_t)r, t) % (uint64_t)cols);
      int pos = 0;
      while (pos < n && c[pos] < x) ++pos;
      if (pos < n && c[pos] == x) continue;
      for (int q = n; q > pos; --q) c[q] = c[q - 1];
      c[pos] = x;
      ++n;
    }
    const int64_t base = r * deg;
    for (int q = 0; q < deg; ++q) {
      col[base + q] = c[q];
      put_val(val, base + q, f64, hash3(seed, 1, (uint64_t)(base + q)), -1.0, 1.0);
    }
  }
}

// ---- banded ------------------------------------------------------------------------------------------
__global__ void banded_kernel(int32_t n, int32_t bw, uint64_t seed, int32_t* rowptr, int32_t* col,
                              void* val, bool f64) {
  const int64_t nnz = (int64_t)n * bw;
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < nnz;
       j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = j / bw, q = j % bw;
    int64_t c0 = r - bw / 2;
    c0 = c0 < 0 ? 0 : (c0 > n - bw ? n - bw : c0);
    col[j] = (int32_t)(c0 + q);
    put_val(val, j, f64, hash3(seed, 1, (uint64_t)j), -1.0, 1.0);
    if (q == 0) rowptr[r] = (int32_t)j;
    if (j == nnz - 1) rowptr[n] = (int32_t)nnz;
  }
}

// ---- R-MAT ---------------------------------------------------------------------------------------------
// Quadrant thresholds on the top 32 bits of the hash: a=.57, a+b=.76, a+b+c=.95 (Graph500).
constexpr uint64_t kRmatA = 2448131358ull;    // floor(0.57 * 2^32)
constexpr uint64_t kRmatAB = 3264175144ull;   // floor(0.76 * 2^32)
constexpr uint64_t kRmatABC = 4080218931ull;  // floor(0.95 * 2^32)
__global__ void rmat_edges_kernel(int32_t scale, int64_t n_edges, uint64_t seed, uint64_t* keys,
                                  uint32_t* w) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n_edges;
       e += (int64_t)gridDim.x * blockDim.x) {
    uint64_t r = 0, c = 0;
    for (int l = 0; l < scale; ++l) {
      const uint64_t t = hash3(seed, 3 + (uint64_t)l, (uint64_t)e) >> 32;
      const int q = t < kRmatA ? 0 : t < kRmatAB ? 1 : t < kRmatABC ? 2 : 3;
      r = (r << 1) | (uint64_t)(q >> 1);
      c = (c << 1) | (uint64_t)(q & 1);
    }
    keys[e] = (r << 32) | c;
    w[e] = (uint32_t)(hash3(seed, 2, (uint64_t)e) >> 40);  // 24-bit numerator of the edge value
  }
}
__global__ void rmat_rowptr_kernel(const uint64_t* __restrict__ keys, int64_t nnz, int32_t rows,
                                   int32_t* rowptr) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r <= rows;
       r += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t target = (uint64_t)r << 32;  // first key of row r
    int64_t lo = 0, hi = nnz;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < target) lo = mid + 1; else hi = mid;
    }
    rowptr[r] = (int32_t)lo;
  }
}
__global__ void rmat_fill_kernel(const uint64_t* __restrict__ keys, const uint64_t* __restrict__ wsum,
                                 int64_t nnz, int32_t* col, void* val, bool f64) {
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < nnz;
       j += (int64_t)gridDim.x * blockDim.x) {
    col[j] = (int32_t)(keys[j] & 0xffffffffull);
    // merged value = (sum of 24-bit numerators) * 2^-24: exact in double, rounded once to float
    const double v = (double)wsum[j] * (1.0 / 16777216.0);
    if (f64) ((double*)val)[j] = v; else ((float*)val)[j] = (float)v;
  }
}

struct RmatScratch {
  int device = -1;
  uint64_t* keys = nullptr;   // unique sorted keys
  uint64_t* wsum = nullptr;
  int64_t nnz = 0;
  int32_t scale = 0, ef = 0;
  uint64_t seed = 0;
};
static RmatScratch g_rmat;

static void rmat_free() {
  if (g_rmat.keys) cudaFree(g_rmat.keys);
  if (g_rmat.wsum) cudaFree(g_rmat.wsum);
  g_rmat = RmatScratch();
}

static int grid_for(int64_t n) { return (int)std::min<int64_t>(ceil_div(std::max<int64_t>(n, 1), 256), kNumSM * 32); }

static int rmat_build(cudaStream_t st, int32_t scale, int32_t ef, uint64_t seed) {
  rmat_free();
  const int64_t ne = (int64_t)ef << scale;
  uint64_t *k_in = nullptr, *k_out = nullptr;
  uint32_t *w_in = nullptr, *w_out = nullptr;
  int64_t* d_num = nullptr;
  void* tmp = nullptr;
  SBLAS_CUDA(cudaMalloc(&k_in, ne * 8));
  SBLAS_CUDA(cudaMalloc(&k_out, ne * 8));
  SBLAS_CUDA(cudaMalloc(&w_in, ne * 4));
  SBLAS_CUDA(cudaMalloc(&w_out, ne * 4));
  rmat_edges_kernel<<<grid_for(ne), 256, 0, st>>>(scale, ne, seed, k_in, w_in);
  SBLAS_LAUNCH_CHECK();
  size_t tb = 0;
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, k_in, k_out, w_in, w_out, ne, 0, 64, st));
  SBLAS_CUDA(cudaMalloc(&tmp, tb));
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, k_in, k_out, w_in, w_out, ne, 0, 64, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  SBLAS_CUDA(cudaFree(tmp));
  tmp = nullptr;
  SBLAS_CUDA(cudaFree(w_in));
  // widen the numerators so the per-key sums cannot overflow, then reduce by key
  SBLAS_CUDA(cudaMalloc(&g_rmat.wsum, ne * 8));
  SBLAS_CUDA(cudaMalloc(&d_num, 8));
  {
    cub::TransformInputIterator<uint64_t, cub::CastOp<uint64_t>, uint32_t*> it(w_out, cub::CastOp<uint64_t>());
    tb = 0;
    SBLAS_CUDA(cub::DeviceReduce::ReduceByKey(nullptr, tb, k_out, k_in, it, g_rmat.wsum, d_num,
                                              cub::Sum(), ne, st));
    SBLAS_CUDA(cudaMalloc(&tmp, tb));
    SBLAS_CUDA(cub::DeviceReduce::ReduceByKey(tmp, tb, k_out, k_in, it, g_rmat.wsum, d_num,
                                              cub::Sum(), ne, st));
  }
  int64_t h_num = 0;
  SBLAS_CUDA(cudaMemcpyAsync(&h_num, d_num, 8, cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  cudaFree(tmp); cudaFree(w_out); cudaFree(k_out); cudaFree(d_num);
  g_rmat.keys = k_in;  // unique keys
  g_rmat.nnz = h_num;
  g_rmat.scale = scale; g_rmat.ef = ef; g_rmat.seed = seed;
  cudaGetDevice(&g_rmat.device);
  return 0;
}

}  // namespace sblas

using namespace sblas;
extern "C" {

int sblas_synth_fill_f64(void* s, double* p, int64_t n, uint64_t seed, uint64_t stream, double lo, double hi) {
  if (n <= 0) return 0;
  fill_kernel<double><<<grid_for(n), 256, 0, (cudaStream_t)s>>>(p, n, seed, stream, lo, hi);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
int sblas_synth_fill_f32(void* s, float* p, int64_t n, uint64_t seed, uint64_t stream, float lo, float hi) {
  if (n <= 0) return 0;
  fill_kernel<float><<<grid_for(n), 256, 0, (cudaStream_t)s>>>(p, n, seed, stream, lo, hi);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
int64_t sblas_synth_laplace2d_nnz(int32_t nx, int32_t ny) {
  return 5ll * nx * ny - 2ll * nx - 2ll * ny;
}
int sblas_synth_laplace2d(void* s, int32_t nx, int32_t ny, int32_t* rowptr, int32_t* col, void* val, int f64) {
  SBLAS_CHECK(nx > 0 && ny > 0 && sblas_synth_laplace2d_nnz(nx, ny) < (1ll << 31), "laplace2d: bad size");
  laplace_kernel<<<grid_for((int64_t)nx * ny + 1), 256, 0, (cudaStream_t)s>>>(nx, ny, rowptr, col, val, f64 != 0);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
int sblas_synth_uniform(void* s, int32_t rows, int32_t cols, int32_t deg, uint64_t seed,
                        int32_t* rowptr, int32_t* col, void* val, int f64) {
  SBLAS_CHECK(rows > 0 && deg > 0 && deg <= kMaxDeg && deg <= cols && (int64_t)rows * deg < (1ll << 31),
              "synth_uniform: bad size (deg <= %d)", kMaxDeg);
  uniform_kernel<<<grid_for((int64_t)rows + 1), 256, 0, (cudaStream_t)s>>>(rows, cols, deg, seed, rowptr, col, val, f64 != 0);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
int sblas_synth_banded(void* s, int32_t n, int32_t bw, uint64_t seed, int32_t* rowptr, int32_t* col,
                       void* val, int f64) {
  SBLAS_CHECK(n > 0 && bw > 0 && bw <= n && (int64_t)n * bw < (1ll << 31), "synth_banded: bad size");
  banded_kernel<<<grid_for((int64_t)n * bw), 256, 0, (cudaStream_t)s>>>(n, bw, seed, rowptr, col, val, f64 != 0);
  SBLAS_LAUNCH_CHECK();
  return 0;
}
int sblas_synth_rmat(void* s, int32_t scale, int32_t ef, uint64_t seed, int64_t* nnz_out,
                     int32_t* rowptr, int32_t* col, void* val, int f64) {
  cudaStream_t st = (cudaStream_t)s;
  SBLAS_CHECK(scale > 0 && scale < 31 && ef > 0 && ((int64_t)ef << scale) < (1ll << 31), "synth_rmat: bad size");
  if (!g_rmat.keys || g_rmat.scale != scale || g_rmat.ef != ef || g_rmat.seed != seed) {
    int rc = rmat_build(st, scale, ef, seed);
    if (rc) return rc;
  }
  if (nnz_out) *nnz_out = g_rmat.nnz;
  if (!col) return 0;
  SBLAS_CHECK(rowptr && val, "synth_rmat: null output");
  const int32_t rows = 1 << scale;
  rmat_rowptr_kernel<<<grid_for((int64_t)rows + 1), 256, 0, st>>>(g_rmat.keys, g_rmat.nnz, rows, rowptr);
  rmat_fill_kernel<<<grid_for(g_rmat.nnz), 256, 0, st>>>(g_rmat.keys, g_rmat.wsum, g_rmat.nnz, col, val, f64 != 0);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaStreamSynchronize(st));
  rmat_free();
  return 0;
}
}
