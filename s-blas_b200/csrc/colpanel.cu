// colpanel.cu — the sliced-ELL device mirror of A that CsrSparseMatrix::sync2gpu builds for SpMV
// ("matrix.h: device CSR layout and padding" is one of the subsystems the north star lets change).
//
// The mirror is a list of P column panels; panel p holds the nonzeros whose column lies in [p*W, (p+1)*W), every row's
// entries in their original order.  Each panel is stored in sliced-ELL form (SELL-32, rows sorted by length inside
// windows of 65 536): 32-row slices stored column-major, so a lane that walks ITS row front to back reads coalesced
// lines with its 31 neighbours, rows above 128 nnz cut into pieces whose raw sums a fix-up kernel adds.  spmv.cu's
// spmv_sell_kernel runs one pass per panel.  Three set-ups, chosen by sblas_plan_build_colpanels_*(n_panels = 0):
//
//   gather-heavy inputs (x larger than 96 MB and consecutive rows far apart, col_jump > cols / 256; the north star's
//     16M x 32 uniform-random shape: x = 134 MB against a 126 MB L2, every gather its own 32-byte DRAM sector, 19 % of
//     the HBM roofline): P >= 2 panels, x slices of <= 48 MB that stay L2-RESIDENT while their panel streams by.
//     y = beta*y + alpha*A*x runs as P back-to-back passes that carry the RAW running row sums in part[]: pass 0
//     writes sum_0, pass p continues the accumulation FROM the carried value (a row is still summed in CSR storage
//     order, one rounded multiply and one rounded add per term: bit-identical to spmv.h:24-30 when the row's columns
//     ascend), the last pass applies alpha and beta.  5.63 -> 2.67 ms.
//   regular inputs of >= 2^20 nnz (padding < 25 %, no row above 4096 nnz): ONE panel — the per-lane kernel over
//     coalesced slices needs no staging, barrier or shared memory and beats the TMA-staged PIPE kernel: banded
//     16M x 32 1.33-1.46 -> 1.08 ms (97 % of the measured HBM copy rate), 2-D Laplacian 76 -> 69-73 us.
//   skewed inputs (R-MAT): one panel in the VIRTUAL-ROW form (build_vsell below): 0.598 -> 0.543 ms.
//
// CSR panels run by the staged PIPE kernel remain as the fallback when the sliced-ELL form pads >= 25 %
// (SBLAS_COLPANEL_SELL=0 forces them).  A mirror is an optimisation: in automatic mode no failure to build one (memory,
// limits) is an error, the plain kernels then take the input.
#include <cub/cub.cuh>

#include "plan.cuh"

namespace sblas {

constexpr int kCpMaxPanels = 32;

struct CpOffsets { int64_t off[kCpMaxPanels]; };

__global__ void vs_iota_kernel(int32_t* p, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = (int32_t)i;
}

// cnt[p*(rows+1) + r] = number of nonzeros of row r in panel p (entry [rows] of every panel = 0)
__global__ void __launch_bounds__(256)
cp_count_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, int32_t rows,
                int32_t W, int32_t P, int32_t* __restrict__ cnt) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r > rows) return;
  int32_t c[kCpMaxPanels];
#pragma unroll
  for (int p = 0; p < kCpMaxPanels; ++p) c[p] = 0;
  if (r < rows) {
    const int32_t s = rowptr[r], e = rowptr[r + 1];
    for (int32_t j = s; j < e; ++j) c[min(colidx[j] / W, P - 1)]++;
  }
  for (int p = 0; p < P; ++p) cnt[(int64_t)p * (rows + 1) + r] = c[p];
}

template <typename T>
__global__ void __launch_bounds__(256)
cp_fill_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, const T* __restrict__ val,
               int32_t rows, int32_t W, int32_t P, const int32_t* __restrict__ prp, CpOffsets offs,
               int32_t* __restrict__ oc, T* __restrict__ ov) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= rows) return;
  int64_t pos[kCpMaxPanels];
  for (int p = 0; p < P; ++p) pos[p] = offs.off[p] + prp[(int64_t)p * (rows + 1) + r];
  const int32_t s = rowptr[r], e = rowptr[r + 1];
  for (int32_t j = s; j < e; ++j) {  // storage order of the row is kept inside every panel
    const int32_t c = colidx[j];
    const int64_t d = pos[min(c / W, P - 1)]++;
    oc[d] = c;
    ov[d] = val[j];
  }
}

// ---- sliced-ELL form of a panel (SELL-C-sigma, Kreutzer et al., C = 32, sigma = 256) --------------------------
// A column panel holds ~8 nnz per row; read as CSR by one lane per row, the lanes of a warp stride through
// col_idx / val by the row length (shared-memory bank conflicts when staged, 8 lines per load when not) and a
// quarter of the threads of a tile idle.  In sliced-ELL form the 32 rows of a slice are stored column-major:
// entry u of the slice's row l sits at sptr[slice] + u*32 + l, so every load of a warp is one coalesced 128 / 256
// byte line, each lane still walks ITS row front to back (CSR order: bit-identical to the CPU loop), there is
// no staging, no barrier, no shared memory (all of it stays L1 for the gathers in flight).  Rows are sorted by
// descending length inside windows of 256 rows (a stable block radix sort) so that a slice is padded only to
// its longest row: ~6 % on binomial row lengths; the sorted position -> row map costs 4 bytes per row and pass.
__global__ void __launch_bounds__(kSellSigma)
sell_sort_kernel(const int32_t* __restrict__ prp, int32_t rows, int32_t* __restrict__ perm, int32_t* __restrict__ w32) {
  using Sort = cub::BlockRadixSort<uint32_t, kSellSigma, 1, int32_t>;
  __shared__ typename Sort::TempStorage tmp;
  const int64_t r = (int64_t)blockIdx.x * kSellSigma + threadIdx.x;
  uint32_t key[1] = {r < rows ? (uint32_t)(prp[r + 1] - prp[r]) : 0u};
  int32_t val[1] = {r < rows ? (int32_t)r : -1};
  Sort(tmp).SortDescending(key, val, 0, 14);   // stable: equal lengths keep row order; lengths <= 4096 + margin
  const int64_t q = (int64_t)blockIdx.x * kSellSigma + threadIdx.x;
  perm[q] = val[0];
  if ((threadIdx.x & (kSellC - 1)) == 0) w32[q / kSellC] = (int32_t)key[0] * kSellC;   // the slice's longest row comes first
}

template <typename T>
__global__ void __launch_bounds__(256)
sell_fill_kernel(const int32_t* __restrict__ prp, const int32_t* __restrict__ pci, const T* __restrict__ pv,
                 const int32_t* __restrict__ perm, const int32_t* __restrict__ sptr, int64_t n_pos,
                 int32_t* __restrict__ oc, T* __restrict__ ov) {
  const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (q >= n_pos) return;
  const int32_t r = perm[q];
  const int64_t s = q / kSellC;
  const int32_t base = sptr[s] + (int32_t)(q % kSellC);
  const int32_t width = (sptr[s + 1] - sptr[s]) / kSellC;
  const int32_t a = r >= 0 ? prp[r] : 0, len = r >= 0 ? prp[r + 1] - a : 0;
  for (int32_t u = 0; u < width; ++u) {
    oc[base + u * kSellC] = u < len ? pci[a + u] : -1;
    ov[base + u * kSellC] = u < len ? pv[a + u] : T(0);
  }
}

// CSR panel p (local row pointers prp, entries pci / pv) -> sliced-ELL; returns the stored entries in sl.padded
template <typename T>
static int build_sell(cudaStream_t st, int32_t rows, const int32_t* prp, const int32_t* pci, const T* pv,
                      ColPanels::Sell* sl) {
  const int32_t n_win = (int32_t)ceil_div(rows, kSellSigma);
  sl->n_slices = n_win * (kSellSigma / kSellC);
  const int64_t n_pos = (int64_t)sl->n_slices * kSellC;
  SBLAS_CUDA(cudaMalloc(&sl->d_perm, sizeof(int32_t) * (size_t)n_pos));
  SBLAS_CUDA(cudaMalloc(&sl->d_sptr, sizeof(int32_t) * ((size_t)sl->n_slices + 1)));
  SBLAS_CUDA(cudaMemsetAsync(sl->d_sptr + sl->n_slices, 0, sizeof(int32_t), st));
  sell_sort_kernel<<<n_win, kSellSigma, 0, st>>>(prp, rows, sl->d_perm, sl->d_sptr);
  SBLAS_LAUNCH_CHECK();
  size_t tb = 0;
  void* tmp = nullptr;
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, sl->d_sptr, sl->d_sptr, sl->n_slices + 1, st));
  SBLAS_CUDA(cudaMalloc(&tmp, tb ? tb : 16));
  cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tb, sl->d_sptr, sl->d_sptr, sl->n_slices + 1, st);
  int32_t total = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&total, sl->d_sptr + sl->n_slices, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(tmp);
  SBLAS_CHECK(e == cudaSuccess, "plan_build_colpanels: sliced-ELL scan: %s", cudaGetErrorString(e));
  sl->padded = total;
  SBLAS_CUDA(cudaMalloc(&sl->d_col, sizeof(int32_t) * (size_t)std::max(total, 4)));
  SBLAS_CUDA(cudaMalloc(&sl->d_val, sizeof(T) * (size_t)std::max(total, 4)));
  sell_fill_kernel<T><<<(int)ceil_div(n_pos, 256), 256, 0, st>>>(prp, pci, pv, sl->d_perm, sl->d_sptr, n_pos, sl->d_col, (T*)sl->d_val);
  SBLAS_LAUNCH_CHECK();
  return 0;
}

// ---- virtual-row sliced-ELL (skewed inputs: R-MAT, config 4) ---------------------------------------------------
// Power-law rows do not fit sliced-ELL as they are (rows of 10^5 nnz beside empty ones), and the merge-path SPLIT
// kernel that takes them is bound by the L1TEX wavefronts of its x gathers.  Here every row longer than
// kVsellPiece = 128 entries is cut into pieces of 128 ("virtual rows"); the virtual rows are sorted by length inside
// windows of 65 536 (one radix sort of 24-bit keys {window, 255 - length}: slices of 32 then hold rows of almost
// equal length, padding < 1 %, while the y rows of a window stay L2-resident), and stored as ONE sliced-ELL panel.
// The per-lane kernel writes whole rows straight to y (the fused epilogue; still summed in storage order by one
// lane) and the raw sums of pieces to d_piece[]; a small fix-up kernel adds the pieces of each cut row (one warp per
// row, pieces in order, shuffle tree) and finishes it.  Measured on config 4: 0.598 -> ~0.5 ms (tools/exp_c4_vrows.py).
__global__ void __launch_bounds__(256)
vs_count_kernel(const int32_t* __restrict__ rowptr, int32_t rows, int32_t* __restrict__ nch1, int32_t* __restrict__ npc) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r > rows) return;
  int32_t a = 0, b = 0;
  if (r < rows) {
    const int32_t len = rowptr[r + 1] - rowptr[r];
    const int32_t n = (len + kVsellPiece - 1) / kVsellPiece;
    a = n > 1 ? n : 1;     // virtual rows of this row (an empty row keeps one empty virtual row: beta*y)
    b = n > 1 ? n : 0;     // piece slots (cut rows only)
  }
  nch1[r] = a;
  npc[r] = b;
}

// per row: its virtual rows' {start, key, info}; cut rows also append {row, first slot, pieces} to the split list
__global__ void __launch_bounds__(256)
vs_expand_kernel(const int32_t* __restrict__ rowptr, int32_t rows, const int32_t* __restrict__ vfirst,
                 const int32_t* __restrict__ pfirst, int32_t* __restrict__ vstart, uint32_t* __restrict__ vkey,
                 int32_t* __restrict__ vinfo, int32_t* __restrict__ split, int32_t* __restrict__ n_split) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int32_t s = rowptr[r], len = rowptr[r + 1] - s;
  const int32_t n = (len + kVsellPiece - 1) / kVsellPiece;
  const int32_t v0 = vfirst[r];
  if (n <= 1) {
    vstart[v0] = s;
    vkey[v0] = ((uint32_t)(v0 / kVsellWindow) << 8) | (uint32_t)(255 - len);
    vinfo[v0] = (int32_t)r;
    return;
  }
  const int32_t p0 = pfirst[r];
  const int32_t k = atomicAdd(n_split, 1);
  split[3 * k] = (int32_t)r; split[3 * k + 1] = p0; split[3 * k + 2] = n;
  for (int32_t c = 0; c < n; ++c) {
    const int32_t v = v0 + c, l = min(kVsellPiece, len - c * kVsellPiece);
    vstart[v] = s + c * kVsellPiece;
    vkey[v] = ((uint32_t)(v / kVsellWindow) << 8) | (uint32_t)(255 - l);
    vinfo[v] = -(p0 + c + 2);
  }
}

// slice widths (x 32) from the sorted keys: the first position of a slice holds its longest virtual row
__global__ void __launch_bounds__(256)
vs_width_kernel(const uint32_t* __restrict__ skey, int64_t nv, int32_t n_slices, int32_t* __restrict__ w32) {
  const int64_t sl = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (sl > n_slices) return;
  int32_t w = 0;
  if (sl < n_slices && sl * kSellC < nv) w = 255 - (int32_t)(skey[sl * kSellC] & 255u);
  w32[sl] = w * kSellC;
}

template <typename T>
__global__ void __launch_bounds__(256)
vs_fill_kernel(const int32_t* __restrict__ ci, const T* __restrict__ val, const uint32_t* __restrict__ skey,
               const int32_t* __restrict__ order, const int32_t* __restrict__ vstart, const int32_t* __restrict__ vinfo,
               const int32_t* __restrict__ sptr, int64_t nv, int64_t n_pos, int32_t* __restrict__ perm,
               int32_t* __restrict__ oc, T* __restrict__ ov) {
  const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (q >= n_pos) return;
  const int64_t sl = q / kSellC;
  const int32_t base = sptr[sl] + (int32_t)(q % kSellC);
  const int32_t width = (sptr[sl + 1] - sptr[sl]) / kSellC;
  int32_t a = 0, len = 0, info = -1;
  if (q < nv) {
    const int32_t v = order[q];
    a = vstart[v];
    len = 255 - (int32_t)(skey[q] & 255u);
    info = vinfo[v];
  }
  perm[q] = info;
  for (int32_t u = 0; u < width; ++u) {
    oc[base + u * kSellC] = u < len ? ci[a + u] : -1;
    ov[base + u * kSellC] = u < len ? val[a + u] : T(0);
  }
}

struct DevBuf {  // frees a temporary on every exit path
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16) == cudaSuccess ? 0 : 1; }
};

// returns 0 and fills *sl (caller frees on failure too); *kept = false when the form pads too much
template <typename T>
static int build_vsell(cudaStream_t st, int32_t rows, const int32_t* rp, int64_t nnz, const int32_t* ci, const T* val,
                       ColPanels::Sell* sl, bool* kept) {
  *kept = false;
  DevBuf nch1, npc, tmp, vstart, vkey, vkey2, vinfo, order_in, order, cnt;
  if (nch1.alloc(sizeof(int32_t) * ((size_t)rows + 1)) || npc.alloc(sizeof(int32_t) * ((size_t)rows + 1)) || cnt.alloc(16)) {
    set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return 1;
  }
  const int rblocks = (int)ceil_div((int64_t)rows + 1, 256);
  vs_count_kernel<<<rblocks, 256, 0, st>>>(rp, rows, (int32_t*)nch1.p, (int32_t*)npc.p);
  SBLAS_LAUNCH_CHECK();
  size_t tb = 0;
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, (int32_t*)nch1.p, (int32_t*)nch1.p, rows + 1, st));
  size_t tb2 = 0;
  if (tmp.alloc(std::max(tb, (size_t)16))) { set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return 1; }
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, (int32_t*)nch1.p, (int32_t*)nch1.p, rows + 1, st));
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, (int32_t*)npc.p, (int32_t*)npc.p, rows + 1, st));
  int32_t h[2] = {0, 0};
  SBLAS_CUDA(cudaMemcpyAsync(&h[0], (int32_t*)nch1.p + rows, 4, cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaMemcpyAsync(&h[1], (int32_t*)npc.p + rows, 4, cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaMemsetAsync(cnt.p, 0, 16, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  const int64_t nv = h[0];
  const int32_t n_piece = h[1];
  SBLAS_CHECK(nv >= rows && nv / kVsellWindow < (1 << 16), "plan_build_colpanels: too many virtual rows");
  const int32_t n_slices = (int32_t)ceil_div(nv, kSellC);
  const int64_t n_pos = (int64_t)n_slices * kSellC;
  const int32_t max_split = (int32_t)std::min<int64_t>(rows, n_piece / 2 + 1);
  if (vstart.alloc(4 * (size_t)nv) || vkey.alloc(4 * (size_t)nv) || vkey2.alloc(4 * (size_t)nv) || vinfo.alloc(4 * (size_t)nv) ||
      order_in.alloc(4 * (size_t)nv) || order.alloc(4 * (size_t)nv)) {
    set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return 1;
  }
  SBLAS_CUDA(cudaMalloc(&sl->d_split, sizeof(int32_t) * 3 * (size_t)std::max(max_split, 1)));
  SBLAS_CUDA(cudaMalloc(&sl->d_piece, sizeof(T) * (size_t)std::max(n_piece, 1)));
  vs_expand_kernel<<<(int)ceil_div(rows, 256), 256, 0, st>>>(rp, rows, (const int32_t*)nch1.p, (const int32_t*)npc.p, (int32_t*)vstart.p,
                                                            (uint32_t*)vkey.p, (int32_t*)vinfo.p, sl->d_split, (int32_t*)cnt.p);
  SBLAS_LAUNCH_CHECK();
  vs_iota_kernel<<<(int)ceil_div(nv, 256), 256, 0, st>>>((int32_t*)order_in.p, nv);
  SBLAS_LAUNCH_CHECK();
  const int key_bits = 8 + std::max(1, 32 - __builtin_clz((unsigned)(nv / kVsellWindow + 1)));
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb2, (const uint32_t*)vkey.p, (uint32_t*)vkey2.p, (const int32_t*)order_in.p,
                                             (int32_t*)order.p, nv, 0, key_bits, st));
  DevBuf tmp2;
  if (tmp2.alloc(tb2)) { set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return 1; }
  SBLAS_CUDA(cub::DeviceRadixSort::SortPairs(tmp2.p, tb2, (const uint32_t*)vkey.p, (uint32_t*)vkey2.p, (const int32_t*)order_in.p,
                                             (int32_t*)order.p, nv, 0, key_bits, st));   // LSD radix sort: stable
  sl->n_slices = n_slices;
  SBLAS_CUDA(cudaMalloc(&sl->d_perm, sizeof(int32_t) * (size_t)n_pos));
  SBLAS_CUDA(cudaMalloc(&sl->d_sptr, sizeof(int32_t) * ((size_t)n_slices + 1)));
  vs_width_kernel<<<(int)ceil_div((int64_t)n_slices + 1, 256), 256, 0, st>>>((const uint32_t*)vkey2.p, nv, n_slices, sl->d_sptr);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, sl->d_sptr, sl->d_sptr, n_slices + 1, st));
  DevBuf tmp3;
  if (tmp3.alloc(tb)) { set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return 1; }
  SBLAS_CUDA(cub::DeviceScan::ExclusiveSum(tmp3.p, tb, sl->d_sptr, sl->d_sptr, n_slices + 1, st));
  int32_t total = 0, n_split = 0;
  SBLAS_CUDA(cudaMemcpyAsync(&total, sl->d_sptr + n_slices, 4, cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaMemcpyAsync(&n_split, cnt.p, 4, cudaMemcpyDeviceToHost, st));
  SBLAS_CUDA(cudaStreamSynchronize(st));
  sl->padded = total;
  sl->n_split = n_split;
  sl->n_piece = n_piece;
  if ((int64_t)total * 4 > nnz * 5 + (int64_t)rows) return 0;   // pads >= 25 %: not kept
  SBLAS_CUDA(cudaMalloc(&sl->d_col, sizeof(int32_t) * (size_t)std::max(total, 4)));
  SBLAS_CUDA(cudaMalloc(&sl->d_val, sizeof(T) * (size_t)std::max(total, 4)));
  vs_fill_kernel<T><<<(int)ceil_div(n_pos, 256), 256, 0, st>>>(ci, val, (const uint32_t*)vkey2.p, (const int32_t*)order.p,
                                                              (const int32_t*)vstart.p, (const int32_t*)vinfo.p, sl->d_sptr, nv, n_pos,
                                                              sl->d_perm, sl->d_col, (T*)sl->d_val);
  SBLAS_LAUNCH_CHECK();
  SBLAS_CUDA(cudaStreamSynchronize(st));   // the temporaries are released on return
  *kept = true;
  return 0;
}

void colpanels_free(sblasPlan_st* plan) {
  ColPanels& cp = plan->cp;
  for (sblasPlan_st* s : cp.sub)
    if (s) sblas_plan_destroy(s);
  if (cp.d_rowptr) cudaFree(cp.d_rowptr);
  if (cp.d_colidx) cudaFree(cp.d_colidx);
  if (cp.d_val) cudaFree(cp.d_val);
  if (cp.d_part) cudaFree(cp.d_part);
  for (ColPanels::Sell& sl : cp.sell) ColPanels::free_sell(sl);
  cp = ColPanels();
}

template <typename T>
static int build_colpanels_n(cudaStream_t st, sblasPlan_st* plan, const int32_t* ci, const T* val, int32_t cols, int P,
                             bool auto_mode);

// rows of a matrix worth a mirror in automatic mode (smaller inputs are launch-bound whatever the kernel)
constexpr int64_t kAutoMirrorMinNnz = 1 << 20;

template <typename T>
static int build_colpanels(cudaStream_t st, sblasPlan_st* plan, const int32_t* ci, const T* val, int32_t cols,
                           int32_t n_panels) {
  SBLAS_CHECK(plan && ci && val && cols > 0, "plan_build_colpanels: bad arguments");
  colpanels_free(plan);
  const int32_t rows = plan->rows;
  if (rows == 0 || plan->nnz == 0) return 0;
  if (n_panels != 0) return build_colpanels_n<T>(st, plan, ci, val, cols, n_panels, false);
  // ---- automatic (what sync2gpu asks for).  SBLAS_SPMV_MIRROR=0 forbids any mirror.
  {
    const char* em = getenv("SBLAS_SPMV_MIRROR");
    if (em && atoi(em) == 0) return 0;
  }
  int P;
  if (measure_gather_heavy(st, plan, plan->d_rowptr, ci, rows, cols, sizeof(T))) {
    // gather-heavy inputs: x slices of SBLAS_COLPANEL_MB (default 48) MB: on the 16M x 32 uniform-random input 3
    // panels (43 MB) run 2.83 ms, 2 (64 MB) 2.98, 4 (32 MB) 2.91, 6 3.21, 8 3.48 (profiles/r02_exp_colpanel_v4_sell.log):
    // fewer passes carry the row sums less often, but the slice must stay L2-resident beside the streaming panel
    static const int mb = [] { const char* e = getenv("SBLAS_COLPANEL_MB"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 48; }();
    P = (int)ceil_div((int64_t)cols * (int64_t)sizeof(T), (int64_t)mb << 20);
    if (P < 2) P = 1;
  } else {
    // every other input of some size whose rows fit the sliced-ELL form (padding < 25 %): ONE panel.  The per-lane
    // walk over coalesced slices needs no staging, no barrier and no shared memory, and measured faster than the
    // TMA-staged PIPE kernel on regular rows: banded 16M x 32 1.33-1.46 -> 1.08 ms (98 % of the measured HBM copy
    // rate), 2-D Laplacian 76.1 -> 73.2 us (profiles/r02_exp_sell_single_panel.log); still one lane per row in storage
    // order, i.e. bit-identical to spmv.h:24-30.  Skewed inputs (R-MAT) fail the padding test and keep SPLIT.
    P = 1;
  }
  if (P == 1 && plan->nnz < kAutoMirrorMinNnz) return 0;
  {  // room for the mirror and its transient CSR copy? (a mirror is an optimisation: never fail sync2gpu for it)
    size_t free_b = 0, total_b = 0;
    const size_t need = (size_t)plan->nnz * (sizeof(T) + 4) * 5 / 2 + (size_t)rows * 64;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < need + (1ull << 30)) { cudaGetLastError(); return 0; }
  }
  if (P > 1 || plan->max_row_nnz <= 4096) {
    if (build_colpanels_n<T>(st, plan, ci, val, cols, P, true) != 0) {
      colpanels_free(plan);
      cudaGetLastError();
      return 0;
    }
    if (plan->cp.n > 0 || P > 1) return 0;
  }
  // skewed rows (the plain form padded >= 25 %, or rows above 4096 nnz): the virtual-row form, one panel.
  // SBLAS_SPMV_VSELL=0 keeps the merge-path SPLIT kernel for such inputs.
  {
    const char* ev = getenv("SBLAS_SPMV_VSELL");
    if (ev && atoi(ev) == 0) return 0;
  }
  if ((int64_t)rows + plan->nnz / kVsellPiece + 1 >= (1ll << 31) - kVsellWindow) return 0;
  ColPanels::Sell sl;
  bool kept = false;
  if (build_vsell<T>(st, rows, plan->d_rowptr, plan->nnz, ci, val, &sl, &kept) != 0 || !kept) {
    ColPanels::free_sell(sl);
    cudaGetLastError();
    return 0;
  }
  ColPanels& cp = plan->cp;
  cp.n = 1;
  cp.panel_cols = (int32_t)(((int64_t)cols + 31) & ~(int64_t)31);
  cp.is_f64 = sizeof(T) == 8;
  cp.sell.assign(1, sl);
  cp.skewed = true;
  return 0;
}

template <typename T>
static int build_colpanels_n(cudaStream_t st, sblasPlan_st* plan, const int32_t* ci, const T* val, int32_t cols, int P,
                             bool auto_mode) {
  const int32_t rows = plan->rows;
  SBLAS_CHECK(P >= 1 && P <= kCpMaxPanels, "plan_build_colpanels: 1..%d panels", kCpMaxPanels);
  const int32_t W = (int32_t)((ceil_div(cols, P) + 31) & ~(int64_t)31);
  ColPanels& cp = plan->cp;
  auto fail = [&](int rc) { colpanels_free(plan); return rc; };
  const size_t rp_elems = (size_t)P * ((size_t)rows + 1);
  if (cudaMalloc(&cp.d_rowptr, rp_elems * sizeof(int32_t)) != cudaSuccess) { set_error("plan_build_colpanels: out of device memory"); cudaGetLastError(); return fail(1); }
  const int blocks = (int)ceil_div((int64_t)rows + 1, 256);
  cp_count_kernel<<<blocks, 256, 0, st>>>(plan->d_rowptr, ci, rows, W, P, cp.d_rowptr);
  ++g_launches;
  // per panel: exclusive scan of the counts in place -> local row pointers; entry [rows] = panel nnz
  size_t tb = 0;
  void* tmp = nullptr;
  if (cub::DeviceScan::ExclusiveSum(nullptr, tb, cp.d_rowptr, cp.d_rowptr, rows + 1, st) != cudaSuccess ||
      cudaMalloc(&tmp, tb ? tb : 16) != cudaSuccess) { set_error("plan_build_colpanels: scan setup failed"); cudaGetLastError(); return fail(1); }
  for (int p = 0; p < P; ++p) {
    int32_t* c = cp.d_rowptr + (size_t)p * ((size_t)rows + 1);
    if (cub::DeviceScan::ExclusiveSum(tmp, tb, c, c, rows + 1, st) != cudaSuccess) { cudaFree(tmp); set_error("plan_build_colpanels: scan failed"); cudaGetLastError(); return fail(1); }
  }
  std::vector<int32_t> tot((size_t)P, 0);
  cudaError_t e = cudaSuccess;
  for (int p = 0; p < P && e == cudaSuccess; ++p)
    e = cudaMemcpyAsync(&tot[(size_t)p], cp.d_rowptr + (size_t)p * ((size_t)rows + 1) + rows, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(tmp);
  if (e != cudaSuccess) { set_error("plan_build_colpanels: %s", cudaGetErrorString(e)); cudaGetLastError(); return fail(1); }
  CpOffsets offs;
  cp.nnz_off.assign((size_t)P + 1, 0);
  int64_t total = 0;
  for (int p = 0; p < P; ++p) {
    cp.nnz_off[(size_t)p] = total;
    offs.off[p] = total;
    total += ((int64_t)tot[(size_t)p] + 3) & ~(int64_t)3;  // every panel starts 16-byte aligned (TMA bulk staging)
  }
  cp.nnz_off[(size_t)P] = total;
  for (int p = P; p < kCpMaxPanels; ++p) offs.off[p] = 0;
  if (cudaMalloc(&cp.d_colidx, (size_t)std::max<int64_t>(total, 4) * sizeof(int32_t)) != cudaSuccess ||
      cudaMalloc(&cp.d_val, (size_t)std::max<int64_t>(total, 4) * sizeof(T)) != cudaSuccess ||
      cudaMalloc(&cp.d_part, (size_t)rows * sizeof(T)) != cudaSuccess) {
    set_error("plan_build_colpanels: out of device memory for the mirror (%lld nnz)", (long long)total);
    cudaGetLastError();
    return fail(1);
  }
  cp_fill_kernel<T><<<(int)ceil_div(rows, 256), 256, 0, st>>>(plan->d_rowptr, ci, val, rows, W, P, cp.d_rowptr, offs,
                                                            cp.d_colidx, (T*)cp.d_val);
  ++g_launches;
  if ((e = cudaGetLastError()) != cudaSuccess) { set_error("plan_build_colpanels: fill launch: %s", cudaGetErrorString(e)); return fail(3); }
  // sub-plans: row statistics + the PIPE tile table of every panel, so that no SpMV call ever builds (and
  // synchronises for) them
  cp.n = P;
  cp.panel_cols = W;
  cp.is_f64 = sizeof(T) == 8;
  cp.sub.assign((size_t)P, nullptr);
  for (int p = 0; p < P; ++p) {
    sblasPlan_st* sub = nullptr;
    if (sblas_plan_create_i32(st, rows, tot[(size_t)p], cp.d_rowptr + (size_t)p * ((size_t)rows + 1), &sub)) return fail(4);
    cp.sub[(size_t)p] = sub;
    {
      // tile of the panel passes: the staged kernel's default (1536 merge items: 2304 and 3456 measured 30-80 %
      // slower — fewer resident CTAs, profiles/r02_exp_colpanel_v2_xrow.log); SBLAS_COLPANEL_TILE overrides
      const char* et = getenv("SBLAS_COLPANEL_TILE");
      const int32_t t = et ? std::max(512, std::min(8192, (atoi(et) + 127) & ~127)) : 0;
      if (t > 0 && 2 * sub->max_row_nnz + 512 <= t) sub->pipe_tile = t;
    }
    bool pipe_ok = false;
    if (prepare_pipe_table(st, sub, &pipe_ok)) return fail(4);
    if (!pipe_ok) {  // a panel row too long for the staging buffers: no mirror, the plain kernels take the input
      colpanels_free(plan);
      return 0;
    }
  }
  // sliced-ELL form of every panel, kept (and the CSR form dropped) when padding stays below 25 %:
  // SBLAS_COLPANEL_SELL=0 keeps the CSR panels and the staged kernel
  const char* es = getenv("SBLAS_COLPANEL_SELL");
  if (!(es && atoi(es) == 0) && (int64_t)P * ((int64_t)rows + kSellSigma) < (1ll << 31)) {
    cp.sell.assign((size_t)P, ColPanels::Sell());
    int64_t padded = 0;
    bool ok = true;
    for (int p = 0; p < P && ok; ++p) {
      const int64_t off = cp.nnz_off[(size_t)p];
      if ((int64_t)tot[(size_t)p] * 5 / 4 + (int64_t)kSellC * rows / 8 >= (1ll << 31)) { ok = false; break; }
      // rows sorted by length in windows of 65 536 by the virtual-row builder (padding < 1 %; rows above 128 nnz cut):
      // against the 256-row block sort below c3v 2.835 -> 2.699 ms, c2 69.8 -> 68.6 us, b32 level
      // (profiles/r02_exp_sell_window_sort.log).  SBLAS_COLPANEL_SORT=0 keeps the 256-row windows.
      const char* esrt = getenv("SBLAS_COLPANEL_SORT");
      if (!(esrt && atoi(esrt) == 0)) {
        bool kept = false;
        if (build_vsell<T>(st, rows, cp.d_rowptr + (size_t)p * ((size_t)rows + 1), tot[(size_t)p], cp.d_colidx + off,
                           (const T*)cp.d_val + off, &cp.sell[(size_t)p], &kept)) return fail(4);
        if (!kept) { ok = false; break; }
      } else
      if (build_sell<T>(st, rows, cp.d_rowptr + (size_t)p * ((size_t)rows + 1), cp.d_colidx + off, (const T*)cp.d_val + off,
                        &cp.sell[(size_t)p])) return fail(4);
      padded += cp.sell[(size_t)p].padded;
    }
    if (ok && padded * 4 <= plan->nnz * 5 + (int64_t)rows) {
      SBLAS_CUDA(cudaStreamSynchronize(st));
      for (sblasPlan_st* sub : cp.sub) sblas_plan_destroy(sub);
      cp.sub.clear();
      cudaFree(cp.d_rowptr); cudaFree(cp.d_colidx); cudaFree(cp.d_val);
      cp.d_rowptr = cp.d_colidx = nullptr;
      cp.d_val = nullptr;
    } else {
      for (ColPanels::Sell& sl : cp.sell) ColPanels::free_sell(sl);
      cp.sell.clear();
    }
  }
  // automatic single panel without the sliced-ELL form = a plain copy of the CSR arrays: pointless, drop it
  if (auto_mode && P == 1 && cp.sell.empty()) colpanels_free(plan);
  return 0;
}

}  // namespace sblas

using namespace sblas;
extern "C" {
int sblas_plan_build_colpanels_f64(void* stream, sblasPlan_t plan, const int32_t* d_colidx, const double* d_val,
                                   int32_t cols, int32_t n_panels) {
  return build_colpanels<double>((cudaStream_t)stream, plan, d_colidx, d_val, cols, n_panels);
}
int sblas_plan_build_colpanels_f32(void* stream, sblasPlan_t plan, const int32_t* d_colidx, const float* d_val,
                                   int32_t cols, int32_t n_panels) {
  return build_colpanels<float>((cudaStream_t)stream, plan, d_colidx, d_val, cols, n_panels);
}
int sblas_plan_colpanel_stats(sblasPlan_t plan, int32_t* n_panels, int32_t* panel_cols) {
  SBLAS_CHECK(plan, "plan_colpanel_stats: null plan");
  if (n_panels) *n_panels = plan->cp.n;
  if (panel_cols) *panel_cols = plan->cp.panel_cols;
  return 0;
}
int sblas_plan_colpanel_format(sblasPlan_t plan, int32_t* is_sell, int64_t* stored_entries) {
  SBLAS_CHECK(plan, "plan_colpanel_format: null plan");
  int64_t n = 0;
  for (const ColPanels::Sell& sl : plan->cp.sell) n += sl.padded;
  if (is_sell) *is_sell = plan->cp.sell.empty() ? 0 : 1;
  if (stored_entries) *stored_entries = plan->cp.sell.empty() ? (plan->cp.n ? plan->nnz : 0) : n;
  return 0;
}
}  // extern "C"
