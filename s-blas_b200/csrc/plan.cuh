// plan.cuh — per-device CSR execution plan: row statistics and nnz-balanced tile tables.
//
// The reference rebuilds cuSPARSE descriptors and work buffers on every call
// (spmv.h:70-101).  Here the equivalent analysis happens once, when the matrix is placed
// on the device (CsrSparseMatrix::sync2gpu, matrix.h:331-397), and is reused by every call.
//
// A tile table partitions the merge list "rows + nnz" (Merrill & Garland's merge path,
// evaluated at row granularity): tile b owns the rows r whose merge coordinate
// f(r) = r*row_weight + rowptr[r] lies in [b*T, (b+1)*T).  f is strictly increasing, so
// the owner of each row is unique, tiles hold at most T/row_weight rows, and at most
// T + (length of their last row) nnz.  Empty-row stretches and dense stretches both cost
// ~T items per tile.
#pragma once
#include <vector>

#include "common.cuh"

struct sblasPlan_st;

namespace sblas {

struct TileTable {
  int32_t tile = 0;        // T, merge items per tile
  int32_t row_weight = 1;  // merge weight of one row
  int32_t n_tiles = 0;
  int32_t* d_tile_row = nullptr;  // [n_tiles + 1], first row of each tile; last = rows
  int32_t* d_tile_nnz = nullptr;  // [n_tiles + 1], rowptr[tile_row[b]]: first nnz of each tile; last = nnz
  int32_t max_tile_nnz = 0;
  int32_t max_tile_rows = 0;
};

// Row-block (R x 1 BCSR) mirror of A for SpMM, see spmm_rb.cu.  Only materialised when
// fill = nnz / (n_blocks * R) >= kRowBlockMinFill.
constexpr double kRowBlockMinFill = 0.5;    // R = 4
constexpr double kRowBlockMinFill8 = 0.7;   // R = 8
constexpr int kRbMaskBits = 0xff;           // presence mask = low byte of colmask.y
struct RowBlocks {
  int32_t R = 0;
  int64_t n_blockrows = 0;
  int64_t n_blocks = 0;          // 0: not materialised
  int32_t max_tile_blocks = 0;   // most blocks in 8 consecutive block rows (tile of the pipelined kernel)
  int32_t max_tile_blocks4 = 0;  // ... in 4 consecutive block rows (2 column slices per CTA)
  int32_t max_tile_blocks2 = 0;  // ... in 2 consecutive block rows (4 column slices per CTA)
  double fill = 0.0;
  bool is_f64 = true;
  int32_t* d_ptr = nullptr;      // [n_blockrows + 1] first block of each block row
  void* d_colmask = nullptr;     // int2 {col, presence mask} per block
  void* d_val = nullptr;         // T[n_blocks * R], 0 where absent
  // tile column lists (tiles of 8 block rows, rows with ascending columns only): the distinct
  // columns of a tile, ascending; block b of the tile carries the position of its column in
  // that list in bits 9.. of colmask[b].y.  The TILEB kernel stages those B rows once per tile.
  int32_t* d_tile_colptr = nullptr;  // [n_tiles8 + 1]
  int32_t* d_tile_cols = nullptr;
  int32_t max_tile_cols = 0;         // 0: no lists (unsorted rows, or not built)
};

// Column-panel mirror of A for gather-heavy SpMV (colpanel.cu): panel p = the nonzeros with a column in
// [p*panel_cols, (p+1)*panel_cols), a CSR of its own; n = 0: not materialised.
struct ColPanels {
  int32_t n = 0;
  int32_t panel_cols = 0;
  bool is_f64 = true;
  int32_t* d_rowptr = nullptr;   // [n][rows + 1] local row pointers of every panel
  int32_t* d_colidx = nullptr;   // panel p at nnz_off[p] (a multiple of 4: 16-byte aligned)
  void* d_val = nullptr;
  void* d_part = nullptr;        // [rows] raw running row sums carried from pass to pass
  std::vector<int64_t> nnz_off;  // [n + 1]
  std::vector<sblasPlan_st*> sub;  // per panel: row statistics + tile tables
  // sliced-ELL form of the panels (SELL-32-sigma256, colpanel.cu): when present the CSR arrays above are
  // released and the passes run on these
  struct Sell {
    int32_t n_slices = 0;
    int64_t padded = 0;           // stored entries (32 x sum of slice widths)
    int32_t* d_perm = nullptr;    // [n_slices * 32] row of every sorted position (-1: none)
    int32_t* d_sptr = nullptr;    // [n_slices + 1] first entry of every slice
    int32_t* d_col = nullptr;     // [padded] column, -1 = padding; entry u of sorted position q at sptr[q/32] + u*32 + q%32
    void* d_val = nullptr;
    // virtual-row form (skewed inputs): rows longer than kVsellPiece entries are cut into pieces; a sorted position
    // then holds a whole row (perm >= 0) or one piece of a cut row (perm = -(slot + 2): its raw sum goes to
    // d_piece[slot]); d_split lists the cut rows as {row, first slot, pieces} for the fix-up kernel
    int32_t n_split = 0;
    int32_t n_piece = 0;
    int32_t* d_split = nullptr;   // [n_split * 3]
    void* d_piece = nullptr;      // T[n_piece]
  };
  static void free_sell(Sell& sl) {
    if (sl.d_perm) cudaFree(sl.d_perm);
    if (sl.d_sptr) cudaFree(sl.d_sptr);
    if (sl.d_col) cudaFree(sl.d_col);
    if (sl.d_val) cudaFree(sl.d_val);
    if (sl.d_split) cudaFree(sl.d_split);
    if (sl.d_piece) cudaFree(sl.d_piece);
    sl = Sell();
  }
  std::vector<Sell> sell;
  bool skewed = false;   // the single virtual-row panel of a skewed input (AUTO keeps SPLIT for it when replicas are written)
};

}  // namespace sblas

struct sblasPlan_st {
  int device = 0;
  int32_t rows = 0;
  int64_t nnz = 0;
  const int32_t* d_rowptr = nullptr;  // not owned
  int32_t max_row_nnz = 0;
  int32_t empty_rows = 0;
  std::vector<sblas::TileTable> tables;
  // scratch for the split (carry) SpMV variant, grown on demand
  void* d_split_scratch = nullptr;
  size_t split_scratch_bytes = 0;
  sblas::RowBlocks rb;
  sblas::ColPanels cp;
  // mean |first column of row r+1 - first column of row r| over a sample of rows (-1 = not measured yet):
  // ~1 for banded / stencil inputs, ~cols/nnz_per_row for random ones.  Measured lazily by the first AUTO SpMV
  // whose x does not fit L2 (spmv.cu, col_jump_kernel).
  double col_jump = -1.0;
  int32_t pipe_tile = 0;  // merge items per tile of the staged SpMV kernel for this plan (0 = the default rule)
};

namespace sblas {

// Drops an ephemeral plan (plan == NULL calls) on every exit path of an op.
struct EphemeralPlan {
  sblasPlan_st* p = nullptr;
  cudaStream_t st = nullptr;
  ~EphemeralPlan() {
    if (p) {
      cudaStreamSynchronize(st);
      sblas_plan_destroy(p);
    }
  }
};

// Residency of a persistent kernel, cached per (kernel, device, block, dynamic shared memory): the
// cudaFuncSetAttribute / cudaOccupancyMaxActiveBlocksPerMultiprocessor / cudaDeviceGetAttribute calls cost
// several microseconds of host time, which is visible beside a 70 us SpMV when made on every call.
int kernel_residency(const void* kern, int block, size_t smem, int* per_sm, int* n_sm);

// Finds or builds the tile table (row_weight, tile) of a plan.  Builds synchronise `stream`.
int plan_get_table(sblasPlan_st* plan, cudaStream_t stream, int32_t row_weight, int32_t tile,
                   const TileTable** out);
int plan_scratch(sblasPlan_st* plan, size_t bytes, void** out);
void rowblocks_free(sblasPlan_st* plan);
void colpanels_free(sblasPlan_st* plan);
// spmv.cu: measures the plan's col_jump statistic when x is larger than ~L2 (one tiny kernel + a 16-byte copy,
// synchronises `st`; cached) and says whether the input is gather-heavy (x does not stay in L2 and
// consecutive rows look at unrelated parts of it)
bool measure_gather_heavy(cudaStream_t st, sblasPlan_st* plan, const int32_t* rp, const int32_t* ci, int32_t rows,
                          int32_t cols, size_t elem_size);
constexpr int kSellC = 32;       // slice height = one warp
constexpr int kSellSigma = 256;  // rows sorted by length inside windows of this many rows (one CTA)
constexpr int kVsellPiece = 128;      // virtual-row form: longest piece of a cut row
constexpr int kVsellWindow = 65536;   // ... virtual rows sorted by length inside windows of this many
// spmv.cu: builds the tile table the PIPE kernel will ask for (so that no SpMV call has to)
int prepare_pipe_table(cudaStream_t st, sblasPlan_st* plan, bool* pipe_ok);

}  // namespace sblas
