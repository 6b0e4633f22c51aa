// colpanel.cu — column-panel device mirror of A for gather-heavy SpMV (x larger than L2).
//
// Why: the north star's SpMV shape (16M rows x 32 nnz/row, uniform-random columns, fp64) has an x of
// 134 MB against a 126 MB L2, and consecutive rows look at unrelated parts of it: every x gather is a
// 32-byte DRAM sector of its own (536.9 M x 32 B = 17 GB of sector traffic beside 6.9 GB of compulsory
// bytes), the kernel is bound by DRAM gathers in flight and sits at 19 % of the HBM roofline.  The lever
// is the cache, and it is a data-layout question ("matrix.h: device CSR layout and padding" is one of
// the subsystems the north star lets change): next to the CSR arrays the reference's API exposes,
// sync2gpu builds a mirror of A split into P COLUMN PANELS, panel p = the nonzeros whose column lies in
// [p*W, (p+1)*W), each panel a CSR of its own (row pointers [rows+1], col_idx / val in the original
// order within a row).  y = beta*y + alpha*A*x then runs as P back-to-back passes of the PIPE kernel, pass p
// touching only x[p*W, (p+1)*W) — a slice sized to stay L2-resident (default <= 48 MB of the 126 MB L2)
// while its panel of A streams by — and carrying the RAW running row sums in a scratch
// vector: pass 0 writes sum_0, pass p continues the accumulation FROM the carried value (so a row
// summed by one lane still adds its products in CSR storage order, one rounded multiply and one
// rounded add each: bit-identical to spmv.h:24-30 when the row's columns ascend), the last pass
// applies alpha and beta.  Extra traffic: P x (row pointers + carried sums read and written) =
// P x 20 B per row, i.e. +19 % on the compulsory bytes at P = 4.
//
// Built automatically (sblas_plan_build_colpanels_*, n_panels = 0) in two cases: gather-heavy inputs (x larger than
// 96 MB and the plan's column-locality statistic says consecutive rows are far apart, col_jump > cols / 256) get
// P >= 2 panels; every other input of >= 1M nnz whose rows fit the sliced-ELL form gets ONE panel, because the
// barrier-free per-lane kernel over coalesced slices also beats the staged kernel on regular rows (see below).
#include <cub/cub.cuh>

#include "plan.cuh"

namespace sblas {

constexpr int kCpMaxPanels = 32;

struct CpOffsets { int64_t off[kCpMaxPanels]; };

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

void colpanels_free(sblasPlan_st* plan) {
  ColPanels& cp = plan->cp;
  for (sblasPlan_st* s : cp.sub)
    if (s) sblas_plan_destroy(s);
  if (cp.d_rowptr) cudaFree(cp.d_rowptr);
  if (cp.d_colidx) cudaFree(cp.d_colidx);
  if (cp.d_val) cudaFree(cp.d_val);
  if (cp.d_part) cudaFree(cp.d_part);
  for (ColPanels::Sell& sl : cp.sell) {
    if (sl.d_perm) cudaFree(sl.d_perm);
    if (sl.d_sptr) cudaFree(sl.d_sptr);
    if (sl.d_col) cudaFree(sl.d_col);
    if (sl.d_val) cudaFree(sl.d_val);
  }
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
  if (P == 1 && (plan->nnz < kAutoMirrorMinNnz || plan->max_row_nnz > 4096)) return 0;
  {  // room for the mirror and its transient CSR copy? (a mirror is an optimisation: never fail sync2gpu for it)
    size_t free_b = 0, total_b = 0;
    const size_t need = (size_t)plan->nnz * (sizeof(T) + 4) * 5 / 2 + (size_t)rows * 32;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < need + (1ull << 30)) { cudaGetLastError(); return 0; }
  }
  if (build_colpanels_n<T>(st, plan, ci, val, cols, P, true) != 0) {
    colpanels_free(plan);
    cudaGetLastError();
    return 0;
  }
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
      for (ColPanels::Sell& sl : cp.sell) {
        if (sl.d_perm) cudaFree(sl.d_perm);
        if (sl.d_sptr) cudaFree(sl.d_sptr);
        if (sl.d_col) cudaFree(sl.d_col);
        if (sl.d_val) cudaFree(sl.d_val);
      }
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
