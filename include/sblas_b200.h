/* include/sblas_b200.h — the C ABI of libsblas_b200.so (B200 / sm_100a).
 *
 * This is the drop-in boundary for the multi-GPU CSR SpMV / SpMM hot path of
 * tartarughina/S-BLAS.  The reference has no FFI layer of its own (header-only C++
 * templates); the entry points below are what a binding for that path
 * would bind, and each one cites the reference interface (file:line in the reference
 * checkout) it replaces.  The C++ drop-in headers next to this file (sblas.h, spmv.h,
 * spmm.h, matrix.h, matrix_um.h, kernel.h, utility.h, config.h) are thin inline
 * wrappers over exactly these symbols.
 *
 * Conventions
 *  - plain pointers and sizes only; `stream` is a cudaStream_t passed as void*
 *    (NULL = the legacy default stream); device pointers belong to the CUDA device that
 *    is current on the calling thread unless the function takes per-GPU pointer arrays;
 *  - every function returns 0 on success, non-zero on failure; sblas_last_error()
 *    returns the message of the last failure on the calling thread;
 *  - there is NO CPU fallback: without a CUDA device every compute entry point fails;
 *  - IdxType is int32 (the only index type the reference's tests instantiate,
 *    unit_test.cu:81, spmv_test.cu:28), DataType is f32 or f64 (utility.h:304-306).
 */
#ifndef SBLAS_B200_H
#define SBLAS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SBLAS_API __attribute__((visibility("default")))
#else
#define SBLAS_API
#endif

#define SBLAS_MAX_GPU 16

/* ---- status ------------------------------------------------------------------ */
SBLAS_API int sblas_version(void);
SBLAS_API const char* sblas_last_error(void);
/* number of visible CUDA devices (0 when there is no driver/GPU) */
SBLAS_API int sblas_device_count(void);
/* number of CUDA kernels this library has launched so far in this process */
SBLAS_API unsigned long long sblas_launch_count(void);

/* NVTX ranges around the phases of the path (the reference marks its load / distribute / tuning phases with
 * nvtxRangePush / Pop in matrix_um.h:435-445, :485-519, :738-741 and its UM drivers): partition + upload in
 * sync2gpu, the tuning hooks, each op, and inside the multi-GPU ops the compute and the finish / stitch-wait phases.
 * No-ops unless a profiler is attached.  Return the nesting level like nvtxRangePushA / nvtxRangePop. */
SBLAS_API int sblas_range_push(const char* name);
SBLAS_API int sblas_range_pop(void);

/* ---- launch options ------------------------------------------------------------ */
enum {
  SBLAS_VARIANT_AUTO = 0,
  SBLAS_SPMV_TILE = 1,    /* nnz-balanced row-aligned merge-path tiles, smem-staged products */
  SBLAS_SPMV_VECTOR = 2,  /* sub-warp (2..32 lanes) per row, no plan needed */
  SBLAS_SPMV_SPLIT = 3,   /* merge-path tiles that split long rows, carry fix-up (skewed inputs) */
  SBLAS_SPMV_PIPE = 4,    /* TILE's tiles walked by persistent CTAs; col_idx/val/rowptr/y of the next
                             tiles staged by TMA bulk copies + LDGSTS behind mbarriers (AUTO's choice) */
  SBLAS_SPMV_COLPANEL = 5, /* one PIPE pass per column panel of the plan's column-panel mirror of A
                              (sblas_plan_build_colpanels_*): each pass gathers from an L2-resident slice of x and
                              carries the raw row sums to the next; AUTO's choice when the plan holds the mirror */
  SBLAS_SPMM_ROWMAJOR = 1, /* internal row-major B/C: one lane group per A row, 128-bit B gathers */
  SBLAS_SPMM_COLMAJOR = 2, /* native col-major B/C (the reference's host layout), no transposes */
  SBLAS_SPMM_ROWBLOCK = 3, /* row-major B/C on the plan's R x 1 row-block mirror of A: one B load feeds
                              up to R adjacent rows (structured inputs); needs
                              sblas_plan_build_rowblocks_* */
  SBLAS_SPMM_ROWBLOCK_PIPE = 4, /* the same on persistent CTAs (producer warp + 8 consumer warps): mirror
                              ranges of the next tiles staged by TMA bulk copies behind mbarriers (AUTO's choice
                              when the plan holds the mirror) */
  SBLAS_SPMM_ROWBLOCK_TILEB = 5 /* ... and the DISTINCT B rows of every tile staged once by TMA too (the
                              mirror's tile column lists; rows with ascending columns, lists that fit
                              shared memory).  On request only: measured level with ROWBLOCK_PIPE */
};
enum {
  SBLAS_FLAG_FMA = 1,          /* SpMV: allow fused multiply-add in the VECTOR variant (the other SpMV kernels
                                  are HBM-bound and always use separately rounded mul/add, bit-identical to
                                  spmv.h:22-31 for one lane per row).  SpMM: fused multiply-add (already the default) */
  SBLAS_FLAG_EXACT = 1024,     /* SpMM: separately rounded multiply and add in CSR order, bit-identical to
                                  sblas_spmm_csr_cpu (spmm.h:56-68), instead of the default fused multiply-add
                                  (which stays within the north star's 1e-12 / 1e-5 relative row tolerance and
                                  halves the FP64 instruction count).  SBLAS_EXACT=1 in the environment or
                                  sblas_set_spmm_exact(1) make it the default; SBLAS_FLAG_FMA then opts out */
  SBLAS_FLAG_STITCH_DIRECT = 2048, /* replicas (sblasOpts.replica) are written by the computing lanes' own st.global
                                  instead of TMA bulk stores of staged row chunks (the default for wide SpMM rows):
                                  needed for NVSwitch multicast addresses, kept as the comparison point */
  SBLAS_FLAG_STITCH_BULK = 4096, /* SBLAS_SPMM_ROWBLOCK_PIPE: replicas written by TMA bulk stores of the 512-byte row
                                  slices parked in the warp's C buffer (measured level with the per-lane stores: opt-in) */
  SBLAS_FLAG_HEAD_PARTIAL = 2, /* local row 0 continues a row begun on the previous GPU
                                  (matrix.h:443-445): its sum goes to carry[0..k), not to y/C */
  SBLAS_FLAG_TAIL_PARTIAL = 4, /* local last row continues on the next GPU: sum -> carry[k..2k) */
  SBLAS_FLAG_NO_L1 = 8,        /* gather x/B with L1::no_allocate */
  SBLAS_FLAG_PREFETCH_B = 16,  /* SBLAS_SPMM_ROWBLOCK_PIPE: the producer warp prefetches the B rows of
                                  the tile it just staged into L1; AUTO sets it for one-slice panels
                                  (k <= 64 fp64 / 128 fp32), where it measured 5% faster */
  SBLAS_FLAG_NO_PREFETCH_B = 32, /* ... never */
  SBLAS_FLAG_SPLIT_SERIAL = 128, /* SBLAS_SPMV_SPLIT: row pointers loaded before the A stream and y read in the
                                   epilogue (the first version; four dependent memory round trips per CTA) */
  SBLAS_FLAG_PREFETCH_C = 512,  /* SBLAS_SPMM_ROWBLOCK_PIPE: consumer warps prefetch the old C rows of the tile two
                                   ahead into L2 */
  SBLAS_FLAG_SPLIT_SEG = 256,   /* SBLAS_SPMV_SPLIT: phase 2 as an in-place segmented sum (row-start bitmap,
                                   reverse segmented scan of the thread leads), balanced for any row mix */
  SBLAS_FLAG_SPLIT_MERGE = 64   /* SBLAS_SPMV_SPLIT: balanced merge-path walk in phase 2 (every thread owns
                                   tile/256 merge items, CTA-wide segmented scan of the carries) instead of a
                                   sub-warp per row; measured slower on config 4 (0.73 against 0.66 ms), opt-in */
};
enum { SBLAS_ROW_MAJOR = 0, SBLAS_COL_MAJOR = 1 }; /* = MajorOrder, matrix.h:37 */

typedef struct sblasOpts {
  int32_t variant;   /* SBLAS_VARIANT_AUTO or one of the enums above */
  int32_t flags;     /* SBLAS_FLAG_* */
  int32_t tile;      /* tile size in merge items, 0 = auto */
  int32_t lanes;     /* lanes per row for the vector variants, 0 = auto */
  int32_t n_replica; /* extra copies of the output to store (peer GPUs' replicas), <= 15 */
  int32_t stages;    /* pipeline depth of SBLAS_SPMV_PIPE (2..4), 0 = auto */
  void* replica[SBLAS_MAX_GPU - 1]; /* same logical base pointer as y / C on each peer */
  void* carry;       /* device buffer of 2*((k+3)&~3) DataType: HEAD sums at [0,k), TAIL sums at
                        [(k+3)&~3, ...) (k = 1 for SpMV) */
} sblasOpts;

/* ---- integer layer: partitioners (host, bit-exact with the reference) ------------- */
/* utility.h:293-300 csr_findRowIdxUsingNnzIdx — row i with rowptr[i] <= idx < rowptr[i+1]
 * (empty rows skipped), -1 if none.  Binary search instead of the reference's O(h) scan. */
SBLAS_API int32_t sblas_csr_find_row_i32(const int32_t* rowptr, int32_t height, int32_t nnz_idx);

/* matrix.h:360-369 CsrSparseMatrix::sync2gpu(segment): nnz split.  avg_nnz uses integer
 * ceil-div (the reference's float32 ceil agrees whenever it is exact).
 * Outputs have n_gpu entries. */
SBLAS_API int sblas_csr_partition_nnz_i32(const int32_t* rowptr, int32_t height, int32_t nnz,
                                          unsigned n_gpu, int32_t* avg_nnz, int32_t* nnz_gpu,
                                          int32_t* starting_row_gpu, int32_t* stoping_row_gpu);
/* matrix.h:370-375: GPU i's local row-pointer array (stop-start+2 entries, :398-404). */
SBLAS_API int sblas_csr_segment_rowptr_i32(const int32_t* rowptr, int32_t avg_nnz, unsigned i_gpu,
                                           int32_t starting_row, int32_t stoping_row,
                                           int32_t nnz_gpu_i, int32_t* out);
/* matrix.h:557-567 DenseMatrix::sync2gpu(segment): split of the leading dimension. */
SBLAS_API int sblas_dense_split_i32(int32_t first_order, unsigned n_gpu, int32_t* avg_val,
                                    int32_t* dim_gpu);
/* nnz-balanced ROW-ALIGNED panels (north star, method 2): panel i = rows
 * [row_begin[i], row_begin[i+1]); row_begin has n_gpu+1 entries.  A row belongs to the
 * panel in which the reference's nnz split puts its first nnz. */
SBLAS_API int sblas_csr_partition_rows_i32(const int32_t* rowptr, int32_t height, unsigned n_gpu,
                                           int32_t* row_begin);

/* ---- Matrix Market loader (mmio_highlevel.h:7-127, :130-281; matrix.h:300-318) ------- */
SBLAS_API int sblas_mtx_info(const char* filename, int32_t* m, int32_t* n, int32_t* nnz,
                             int32_t* is_symmetric);
SBLAS_API int sblas_mtx_data(const char* filename, int32_t* rowptr, int32_t* colidx, double* val);
/* host threads of the loader: 0 = automatic (hardware threads, >= 256 KB of text each; the
 * environment variable SBLAS_MTX_THREADS overrides), 1 = serial.  The result does not depend on it. */
SBLAS_API void sblas_mtx_set_threads(int n);
/* The coordinate entries as the file holds them (no symmetric expansion), 0-based, file order: what
 * CooSparseMatrix(const char*) reads (matrix.h:142-176).  pattern files give 1.0 (the reference's
 * "%d %d %lg" would misread them); nz = the announced entry count. */
SBLAS_API int sblas_mtx_coo_info(const char* filename, int32_t* m, int32_t* n, int32_t* nz);
SBLAS_API int sblas_mtx_coo_data(const char* filename, int32_t* rowidx, int32_t* colidx, double* val);

/* ---- format containers beside CSR (matrix.h:40-272, :457-506), on the device ----
 * CsrToCsc (matrix.h:40-64): stable by column — inside a column the entries keep CSR order —
 * bit-identical to the reference's CPU loop.  CscToCsr (matrix.h:67-91) is the mirror image; note
 * the reference's argument order (n, m, ...).  All pointers are device pointers on the stream's
 * device; the call returns after the conversion finished (temporaries are released). */
SBLAS_API int sblas_csr2csc_f64(void* stream, int32_t m, int32_t n, int64_t nnz, const int32_t* d_rowptr,
                                const int32_t* d_colidx, const double* d_val, int32_t* d_cscColPtr,
                                int32_t* d_cscRowIdx, double* d_cscVal);
SBLAS_API int sblas_csr2csc_f32(void* stream, int32_t m, int32_t n, int64_t nnz, const int32_t* d_rowptr,
                                const int32_t* d_colidx, const float* d_val, int32_t* d_cscColPtr,
                                int32_t* d_cscRowIdx, float* d_cscVal);
SBLAS_API int sblas_csc2csr_f64(void* stream, int32_t n, int32_t m, int64_t nnz, const int32_t* d_colptr,
                                const int32_t* d_rowidx, const double* d_val, int32_t* d_csrRowPtr,
                                int32_t* d_csrColIdx, double* d_csrVal);
SBLAS_API int sblas_csc2csr_f32(void* stream, int32_t n, int32_t m, int64_t nnz, const int32_t* d_colptr,
                                const int32_t* d_rowidx, const float* d_val, int32_t* d_csrRowPtr,
                                int32_t* d_csrColIdx, float* d_csrVal);
/* CooSparseMatrix::sortByRow (matrix.h:221-237): in-place sort by (row, col); equal pairs keep their
 * order (the reference's qsort leaves that open). */
SBLAS_API int sblas_coo_sort_f64(void* stream, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col,
                                 double* d_val);
SBLAS_API int sblas_coo_sort_f32(void* stream, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col,
                                 float* d_val);

/* ---- per-device CSR plan (tile tables + row statistics; built once at sync2gpu) ------ */
typedef struct sblasPlan_st* sblasPlan_t;
typedef struct sblasCsrStats {
  int64_t nnz;
  int32_t rows;
  int32_t max_row_nnz;
  int32_t empty_rows;
  int32_t spmv_variant; /* what AUTO resolves to */
  int32_t spmv_tile, spmv_tiles, spmv_smem_bytes;
  int32_t spmm_tile, spmm_tiles;
} sblasCsrStats;
SBLAS_API int sblas_plan_create_i32(void* stream, int32_t rows, int64_t nnz,
                                    const int32_t* d_rowptr, sblasPlan_t* plan);
SBLAS_API int sblas_plan_stats(sblasPlan_t plan, sblasCsrStats* out);
/* Column-locality statistic of the plan: mean |first column of row r+1 - first column of row r| over a row
 * sample; -1 until an AUTO SpMV with an x larger than L2 measured it.  AUTO takes the sub-warp-per-row kernel
 * instead of the staged one when it exceeds cols / 256 (uniform-random columns: every gather its own sector). */
SBLAS_API int sblas_plan_col_jump(sblasPlan_t plan, double* out);
/* Takes at plan time what the first AUTO SpMV on the plan would otherwise do inside the call (and synchronise
 * the stream for): the column-locality statistic when x (cols * elem_size bytes) is larger than ~L2, and the tile
 * table of the staged kernel.  After it, sblas_spmv_csr_* on this plan never synchronises or allocates.
 * CsrSparseMatrix::sync2gpu calls it (matrix.h:331-397 is where the reference does its per-GPU setup). */
SBLAS_API int sblas_plan_prepare_i32(void* stream, sblasPlan_t plan, const int32_t* d_colidx, int32_t cols,
                                     int32_t elem_size);
/* Column-panel mirror of A for gather-heavy SpMV ("matrix.h device CSR layout", built inside
 * CsrSparseMatrix::sync2gpu next to the CSR arrays of matrix.h:376-387): A split into n_panels column panels,
 * each a CSR of its own that keeps the storage order of every row.  SpMV then runs one pass per panel, pass p
 * gathering only from an L2-resident slice of x, the raw running row sums carried from pass to pass (rows whose
 * columns ascend are still summed in CSR order).  n_panels = 0: automatic — x slices of at most 48 MB (environment
 * SBLAS_COLPANEL_MB) when x (cols elements) exceeds 96 MB and the plan's col_jump statistic exceeds cols / 256
 * (uniform-random columns); else ONE panel when the matrix has >= 2^20 nnz, no row above 4096 nnz and its sliced-ELL
 * form pads by < 25 % (the per-lane kernel over coalesced 32-row slices is also the fastest SpMV on regular rows);
 * nothing for skewed or small inputs, when device memory is short, or with SBLAS_SPMV_MIRROR=0 in the environment.
 * n_panels >= 1 forces that many panels (<= 32).  The mirror holds a COPY of col_idx / val: rebuild it after
 * changing them. */
SBLAS_API int sblas_plan_build_colpanels_f64(void* stream, sblasPlan_t plan, const int32_t* d_colidx,
                                             const double* d_val, int32_t cols, int32_t n_panels);
SBLAS_API int sblas_plan_build_colpanels_f32(void* stream, sblasPlan_t plan, const int32_t* d_colidx,
                                             const float* d_val, int32_t cols, int32_t n_panels);
SBLAS_API int sblas_plan_colpanel_stats(sblasPlan_t plan, int32_t* n_panels, int32_t* panel_cols);
/* Storage of the panels: *is_sell = 1 when they are kept in sliced-ELL form (32-row slices stored column-major,
 * rows sorted by length inside windows of 256 rows; taken when its padding stays below 25 %, environment
 * SBLAS_COLPANEL_SELL=0 forbids it), 0 for CSR panels run by the staged kernel; *stored_entries = entries held
 * including padding. */
SBLAS_API int sblas_plan_colpanel_format(sblasPlan_t plan, int32_t* is_sell, int64_t* stored_entries);
/* Row-block mirror of A for SpMM ("matrix.h device CSR layout", built inside
 * CsrSparseMatrix::sync2gpu next to the CSR arrays of matrix.h:376-387): R (= 4) consecutive rows
 * form a block row, every distinct column of a block row becomes one block {col, presence mask,
 * R values}.  The arrays are only kept when fill = nnz / (blocks * R) >= 0.5; otherwise only the
 * statistic is recorded and SpMM stays on the CSR kernel.  The mirror holds a COPY of the values:
 * rebuild it after changing d_val.  d_colidx / d_val: the same device arrays later passed to
 * sblas_spmm_csr_*.  R = 0 picks the default; a negative R keeps the mirror whatever the fill. */
SBLAS_API int sblas_plan_build_rowblocks_f64(void* stream, sblasPlan_t plan, const int32_t* d_colidx,
                                             const double* d_val, int32_t R);
SBLAS_API int sblas_plan_build_rowblocks_f32(void* stream, sblasPlan_t plan, const int32_t* d_colidx,
                                             const float* d_val, int32_t R);
SBLAS_API int sblas_plan_rowblock_stats(sblasPlan_t plan, int32_t* R, int64_t* n_blocks, double* fill);
SBLAS_API int sblas_plan_destroy(sblasPlan_t plan);

/* ---- single-device kernels ---------------------------------------------------------- */
/* y <- beta*y + alpha*A*x on the current device.  Replaces cusparseSpMV at spmv.h:96-106
 * plus the axpby pass of spmv.h:138 / kernel.h:28-37 (fused into the epilogue).
 * plan may be NULL (an internal plan is then built and dropped on every call). */
SBLAS_API int sblas_spmv_csr_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz,
                                 const int32_t* d_rowptr, const int32_t* d_colidx,
                                 const double* d_val, const double* d_x, double* d_y, double alpha,
                                 double beta, sblasPlan_t plan, const sblasOpts* opts);
SBLAS_API int sblas_spmv_csr_f32(void* stream, int32_t rows, int32_t cols, int64_t nnz,
                                 const int32_t* d_rowptr, const int32_t* d_colidx,
                                 const float* d_val, const float* d_x, float* d_y, float alpha,
                                 float beta, sblasPlan_t plan, const sblasOpts* opts);

/* y <- beta*y + alpha*A*x on the current device WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same result contract as sblas_spmv_csr_f64 (the merge-path SPLIT
 * kernels: rows of any length, reduction order differs from the CPU loop within the 1e-12 row tolerance); its
 * scratch (split table + per-tile carries) lives in a fixed 3.25 MiB device array inside the library, so
 *   - rows + nnz must not exceed 201 326 592 (131 072 tiles of 1536 merge items) — larger inputs are rejected
 *     with a clear error and nothing is launched;
 *   - the call makes no cudaMalloc / cudaFree / allocating library call, takes no plan and synchronises nothing
 *     (three kernel launches on `stream`);
 *   - every call on a device shares the array: calls that OVERLAP on the GPU (different streams or host threads)
 *     race — serialise them; nothing needs clearing between calls (each word read was written in the same call). */
SBLAS_API int sblas_spmv_csr_noalloc_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz,
                                         const int32_t* d_rowptr, const int32_t* d_colidx, const double* d_val,
                                         const double* d_x, double* d_y, double alpha, double beta);

/* C <- beta*C + alpha*A*B on the current device, k dense columns.  Replaces cusparseSpMM
 * at spmm.h:135-149 (method 1) / :240-251 (method 2) plus the axpby of spmm.h:283.
 * layout = SBLAS_ROW_MAJOR: B[c*ldb + n], C[r*ldc + n] (the internal device layout);
 * layout = SBLAS_COL_MAJOR: B[n*ldb + c], C[n*ldc + r] (the reference's layout, spmm.h:121-132). */
SBLAS_API int sblas_spmm_csr_f64(void* stream, int32_t rows, int32_t cols, int64_t nnz,
                                 const int32_t* d_rowptr, const int32_t* d_colidx,
                                 const double* d_val, int32_t k, int32_t layout, const double* d_B,
                                 int64_t ldb, double* d_C, int64_t ldc, double alpha, double beta,
                                 sblasPlan_t plan, const sblasOpts* opts);
SBLAS_API int sblas_spmm_csr_f32(void* stream, int32_t rows, int32_t cols, int64_t nnz,
                                 const int32_t* d_rowptr, const int32_t* d_colidx,
                                 const float* d_val, int32_t k, int32_t layout, const float* d_B,
                                 int64_t ldb, float* d_C, int64_t ldc, float alpha, float beta,
                                 sblasPlan_t plan, const sblasOpts* opts);

/* Process-wide default arithmetic of sblas_spmm_csr_* calls that carry neither SBLAS_FLAG_EXACT nor
 * SBLAS_FLAG_FMA: on != 0 -> separately rounded mul/add (bit-identical to spmm.h:56-68), 0 -> fused
 * multiply-add (the initial default unless SBLAS_EXACT=1 is in the environment).  Returns the previous
 * setting.  For callers that reach the kernels through the C++ headers and cannot pass flags. */
SBLAS_API int sblas_set_spmm_exact(int on);

/* kernel.h:28-37 denseVector_plusEqual_denseVector: v0[i] = v0[i]*beta + v1[i]*alpha. */
SBLAS_API int sblas_axpby_f64(void* stream, double* d_v0, const double* d_v1, double alpha,
                              double beta, int64_t n);
SBLAS_API int sblas_axpby_f32(void* stream, float* d_v0, const float* d_v1, float alpha,
                              float beta, int64_t n);
/* kernel.h:19-25 denseVector_plusEqual_scalar (with the index bug at :22 fixed):
 * v[i] = v[i]*beta + val. */
SBLAS_API int sblas_scal_add_f64(void* stream, double* d_v, double val, double beta, int64_t n);
/* calibration: grid*256 threads each run 8 chains of `iters` fp64 multiply-adds (mode 0: DFMA,
   mode 1: separately rounded DMUL + DADD as the bit-exact SpMM path issues them); time it with
   events to get the FP64 roof beside the HBM one */
SBLAS_API int sblas_fp64_rate(void* stream, int32_t mode, int32_t grid, int32_t iters, double* d_out);
SBLAS_API int sblas_scal_add_f32(void* stream, float* d_v, float val, float beta, int64_t n);

/* out[b*out_ld + a] = in[a*in_ld + b], a < n_major, b < n_minor.  Converts the reference's
 * col-major dense operands (matrix.h:37) to the internal row-major device layout and back
 * (inside DenseMatrix::sync2gpu / sync2cpu, matrix.h:541-570, :595-612). */
SBLAS_API int sblas_transpose_f64(void* stream, const double* d_in, int64_t in_ld, int64_t n_major,
                                  int64_t n_minor, double* d_out, int64_t out_ld);
SBLAS_API int sblas_transpose_f32(void* stream, const float* d_in, int64_t in_ld, int64_t n_major,
                                  int64_t n_minor, float* d_out, int64_t out_ld);

/* ---- memory plumbing used by the C++ drop-in headers -------------------------------------
 * Replaces the allocation macros of utility.h:86-127 (SAFE_ALOC_HOST = cudaMallocHost,
 * SAFE_ALOC_GPU = cudaMalloc, SAFE_FREE_*) and the cudaMemcpy calls of matrix.h:341-387,
 * :541-570, :595-612, :688-713.  `gpu` is a LOGICAL GPU ordinal (sblas_device_of).  With these
 * the headers need no CUDA toolkit on the include path.
 * sblas_host_alloc returns pinned memory; with no CUDA driver at all it returns ordinary
 * aligned memory so that the host-side integer layer can still be exercised (no kernel can
 * run in that case anyway). */
SBLAS_API int sblas_host_alloc(size_t bytes, void** h_ptr);
SBLAS_API int sblas_host_free(void* h_ptr);
SBLAS_API int sblas_dev_alloc(int gpu, size_t bytes, void** d_ptr);
SBLAS_API int sblas_dev_free(int gpu, void* d_ptr);
SBLAS_API int sblas_copy_h2d(int gpu, void* d_dst, const void* h_src, size_t bytes);
SBLAS_API int sblas_copy_d2h(int gpu, void* h_dst, const void* d_src, size_t bytes);
SBLAS_API int sblas_copy_d2d(int gpu, void* d_dst, const void* d_src, size_t bytes);
SBLAS_API int sblas_dev_memset(int gpu, void* d_ptr, int value, size_t bytes);
SBLAS_API int sblas_dev_sync(int gpu);
/* sblas_plan_create_i32 and sblas_axpby_* (kernel.h:28-37 as launched by
 * plusDenseVectorGPU / plusDenseMatrixGPU, matrix.h:613-625, :714-726) on a logical GPU's
 * default stream; alpha/beta are narrowed to float when is_f64 == 0. */
SBLAS_API int sblas_plan_create_on(int gpu, int32_t rows, int64_t nnz, const int32_t* d_rowptr,
                                   sblasPlan_t* plan);
SBLAS_API int sblas_axpby_on(int gpu, int is_f64, void* d_v0, const void* d_v1, double alpha,
                             double beta, int64_t n);
SBLAS_API int sblas_plan_build_rowblocks_on(int gpu, int is_f64, sblasPlan_t plan, const int32_t* d_colidx,
                                            const void* d_val);
/* sblas_plan_prepare_i32 + sblas_plan_build_colpanels_* (automatic) on a logical GPU's default stream */
SBLAS_API int sblas_plan_prepare_spmv_on(int gpu, int is_f64, sblasPlan_t plan, const int32_t* d_colidx,
                                         const void* d_val, int32_t cols);
SBLAS_API int sblas_scal_add_on(int gpu, int is_f64, void* d_v, double val, double beta, int64_t n);
/* CsrToCsc / CscToCsr (matrix.h:40-91) and CooSparseMatrix::sortByRow (matrix.h:221-237) on a logical
 * GPU's default stream.  (m, n, d_ptr, d_idx) = (rows, cols, rowptr, colidx) for CSR -> CSC and
 * (cols, rows, colptr, rowidx) for CSC -> CSR. */
SBLAS_API int sblas_compressed_transpose_on(int gpu, int is_f64, int32_t m, int32_t n, int64_t nnz,
                                            const int32_t* d_ptr, const int32_t* d_idx, const void* d_val,
                                            int32_t* d_out_ptr, int32_t* d_out_idx, void* d_out_val);
SBLAS_API int sblas_coo_sort_on(int gpu, int is_f64, int64_t nnz, int32_t height, int32_t* d_row, int32_t* d_col,
                                void* d_val);
/* DenseMatrix::sync2gpu / sync2cpu for one GPU's slice (matrix.h:541-570, :595-612).  h is the
 * host slice: `dim` leading-dimension lines of `second` contiguous elements each (for the
 * col_major operands of SpMM: dim columns of `second` rows).  dev_layout SBLAS_COL_MAJOR keeps
 * that layout on the device (ld = second), SBLAS_ROW_MAJOR stores the transposed slice
 * [second x ld] with ld = dim rounded up to 16 bytes (the layout the ROWMAJOR kernel wants).
 * The upload allocates *d_out (release with sblas_dev_free). */
SBLAS_API int sblas_dense_upload_f64(int gpu, const double* h, int64_t second, int64_t dim,
                                     int32_t dev_layout, double** d_out, int64_t* ld_out);
SBLAS_API int sblas_dense_upload_f32(int gpu, const float* h, int64_t second, int64_t dim,
                                     int32_t dev_layout, float** d_out, int64_t* ld_out);
SBLAS_API int sblas_dense_download_f64(int gpu, const double* d, int64_t ld, int64_t second,
                                       int64_t dim, int32_t dev_layout, double* h);
SBLAS_API int sblas_dense_download_f32(int gpu, const float* d, int64_t ld, int64_t second,
                                       int64_t dim, int32_t dev_layout, float* h);
/* DenseMatrix::sync2cpu of every GPU's slice in one call (sblas_spmm_csr_v1 downloads all the C slices inside the
 * op, spmm.h:151 -> matrix.h:599-606): slice i (dim[i] lines of `second` elements, device pointer d[i] on logical
 * GPU i, pitch ld[i]) -> host pointer h[i].  All downloads are enqueued before any is waited for, so the GPUs' host
 * links work in parallel; returns when every slice is on the host.  h should be pinned (sblas_host_alloc). */
SBLAS_API int sblas_dense_download_multi_f64(unsigned n_gpu, const double* const* d, const int64_t* ld, int64_t second,
                                             const int64_t* dim, int32_t dev_layout, double* const* h);
SBLAS_API int sblas_dense_download_multi_f32(unsigned n_gpu, const float* const* d, const int64_t* ld, int64_t second,
                                             const int64_t* dim, int32_t dev_layout, float* const* h);
/* The same two moves for a caller that re-sends a dense operand on every step (what a solver loop around
 * sblas_spmm_csr_v2 does with B.sync2gpu / C.sync2cpu, spmm_test.cu:68-84), streamed and allocation-free:
 * rows [r0, r0 + nr) of the col-major HOST matrix h (element (r, n) at h[n*h_ld + r], k columns; pinned memory
 * for real overlap) <-> the row-major DEVICE block d (element (r - r0, n) at d[(r - r0)*ld + n]) on the current
 * device.  Work is enqueued on `stream`: the PCIe copy of one row chunk overlaps the layout conversion of the
 * previous one (a per-device helper stream per direction, so uploads and downloads overlap too), nothing is
 * allocated per call and nothing synchronises; after a download, `stream` completes only when the bytes are on
 * the host. */
SBLAS_API int sblas_dense_upload_rows_f64(void* stream, const double* h, int64_t h_ld, int64_t r0, int64_t nr,
                                          int32_t k, double* d_dst, int64_t ld);
SBLAS_API int sblas_dense_upload_rows_f32(void* stream, const float* h, int64_t h_ld, int64_t r0, int64_t nr,
                                          int32_t k, float* d_dst, int64_t ld);
SBLAS_API int sblas_dense_download_rows_f64(void* stream, const double* d_src, int64_t ld, int64_t nr, int32_t k,
                                            double* h, int64_t h_ld, int64_t r0);
SBLAS_API int sblas_dense_download_rows_f32(void* stream, const float* d_src, int64_t ld, int64_t nr, int32_t k,
                                            float* h, int64_t h_ld, int64_t r0);
/* Managed memory of the matrix_um.h variants: cudaMallocManaged (matrix_um.h:60-75),
 * cudaMemAdvise / cudaMemPrefetchAsync (matrix_um.h:529-614).  gpu = -1 means the CPU. */
enum {
  SBLAS_ADVISE_SET_READ_MOSTLY = 1,
  SBLAS_ADVISE_UNSET_READ_MOSTLY = 2,
  SBLAS_ADVISE_SET_PREFERRED_LOCATION = 3,
  SBLAS_ADVISE_UNSET_PREFERRED_LOCATION = 4,
  SBLAS_ADVISE_SET_ACCESSED_BY = 5,
  SBLAS_ADVISE_UNSET_ACCESSED_BY = 6
};
SBLAS_API int sblas_managed_alloc(size_t bytes, void** ptr);
SBLAS_API int sblas_managed_free(void* ptr);
SBLAS_API int sblas_mem_advise(void* ptr, size_t bytes, int advice, int gpu);
SBLAS_API int sblas_mem_prefetch(void* ptr, size_t bytes, int gpu);

/* ---- multi-GPU ops, one process driving GPUs 0..n_gpu-1 (the reference's execution model:
 *      spmv.h:52-55, spmm.h:101-104; streams/events/peer access live in a cached context) ---- */

/* Physical CUDA device of logical GPU i.  Identity (the reference's cudaSetDevice(i_gpu),
 * spmv.h:55) unless the environment variable SBLAS_DEVICE_MAP="d0,d1,..." remaps it, which
 * lets an n-way partition run on fewer devices (several logical GPUs per device). */
SBLAS_API int sblas_device_of(int logical_gpu);

/* sblas_spmv_csr_v1, spmv.h:36-139.  A nnz-segmented exactly as matrix.h:356-394 leaves it
 * (per-GPU local rowptr / col / val slices and the three partition tables); x and y
 * replicated.  On return EVERY replica y_gpu[i] holds beta*y + alpha*A*x (the reference's
 * post-condition, matrix.h:714-726).  Owner-computes rows + direct peer stores from the
 * kernel epilogue replace the reference's zero-buffer + ncclAllReduce + axpby passes; rows
 * shared by neighbouring GPUs are finished by a carry fix-up.  plans[i] may be NULL. */
SBLAS_API int sblas_mgpu_spmv_csr_v1_f64(unsigned n_gpu, int32_t height, int32_t width,
                                         const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const double* const* val_gpu, const int32_t* nnz_gpu,
                                         const int32_t* starting_row_gpu,
                                         const int32_t* stoping_row_gpu,
                                         const double* const* x_gpu, double* const* y_gpu,
                                         double alpha, double beta, const sblasPlan_t* plans);
SBLAS_API int sblas_mgpu_spmv_csr_v1_f32(unsigned n_gpu, int32_t height, int32_t width,
                                         const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const float* const* val_gpu, const int32_t* nnz_gpu,
                                         const int32_t* starting_row_gpu,
                                         const int32_t* stoping_row_gpu,
                                         const float* const* x_gpu, float* const* y_gpu,
                                         float alpha, float beta, const sblasPlan_t* plans);

/* sblas_spmm_csr_v1 (method 1), spmm.h:84-161: A replicated, B and C split by columns
 * (dim_gpu[i] columns on GPU i, matrix.h:557-567).  No inter-GPU traffic.  layout/ld as in
 * sblas_spmm_csr_*; ldb/ldc arrays have one entry per GPU. */
SBLAS_API int sblas_mgpu_spmm_csr_v1_f64(unsigned n_gpu, int32_t height, int32_t width,
                                         int32_t nnz, const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const double* const* val_gpu, const int32_t* dim_gpu,
                                         int32_t layout, const double* const* B_gpu,
                                         const int64_t* ldb, double* const* C_gpu,
                                         const int64_t* ldc, double alpha, double beta,
                                         const sblasPlan_t* plans);
SBLAS_API int sblas_mgpu_spmm_csr_v1_f32(unsigned n_gpu, int32_t height, int32_t width,
                                         int32_t nnz, const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const float* const* val_gpu, const int32_t* dim_gpu,
                                         int32_t layout, const float* const* B_gpu,
                                         const int64_t* ldb, float* const* C_gpu,
                                         const int64_t* ldc, float alpha, float beta,
                                         const sblasPlan_t* plans);

/* sblas_spmm_csr_v2 (method 2), spmm.h:164-284: A nnz-segmented, B and C replicated.  On
 * return every replica C_gpu[i] holds beta*C + alpha*A*B.  Same owner-computes + peer-store
 * stitch as the SpMV above (replaces C_copy + ncclAllReduce(h*k) + axpby, spmm.h:182,260,283). */
SBLAS_API int sblas_mgpu_spmm_csr_v2_f64(unsigned n_gpu, int32_t height, int32_t width,
                                         const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const double* const* val_gpu, const int32_t* nnz_gpu,
                                         const int32_t* starting_row_gpu,
                                         const int32_t* stoping_row_gpu, int32_t k, int32_t layout,
                                         const double* const* B_gpu, int64_t ldb,
                                         double* const* C_gpu, int64_t ldc, double alpha,
                                         double beta, const sblasPlan_t* plans);
SBLAS_API int sblas_mgpu_spmm_csr_v2_f32(unsigned n_gpu, int32_t height, int32_t width,
                                         const int32_t* const* rowptr_gpu,
                                         const int32_t* const* colidx_gpu,
                                         const float* const* val_gpu, const int32_t* nnz_gpu,
                                         const int32_t* starting_row_gpu,
                                         const int32_t* stoping_row_gpu, int32_t k, int32_t layout,
                                         const float* const* B_gpu, int64_t ldb,
                                         float* const* C_gpu, int64_t ldc, float alpha,
                                         float beta, const sblasPlan_t* plans);

/* One process per GPU (torch.distributed / MPI style launch): replicas that every rank of the
 * node can store into.  ipc_alloc = cudaMalloc + cudaIpcGetMemHandle (64-byte handle to ship to
 * the peers), ipc_open maps a peer's buffer; the mapped pointers go into sblasOpts.replica so
 * the SpMV/SpMM epilogue writes every finished row into all replicas over NVLink. */
SBLAS_API int sblas_ipc_alloc(size_t bytes, void** d_ptr, unsigned char handle[64]);
SBLAS_API int sblas_ipc_open(const unsigned char handle[64], void** d_ptr);
SBLAS_API int sblas_ipc_close(void* d_ptr);
SBLAS_API int sblas_ipc_free(void* d_ptr);

/* Device-side barrier between the ranks of one node (one process per GPU), enqueued on `stream`:
 * rank r stores `epoch` into slot r of every peer's flag array (d_peer_flags[t] = rank t's array
 * of >= world int32, zero-initialised, mapped with sblas_ipc_open; entry [rank] = the local array)
 * and waits until all slots of its own array reach `epoch`.  Because the flag stores follow the
 * SpMV/SpMM kernel in stream order, the wait returning means every peer's rows are in the local
 * replica: this completes the peer-store stitch with no host synchronisation and no collective
 * (the reference needs ncclAllReduce + two OpenMP barriers here, spmv.h:108-120).  Epochs must
 * increase by 1 per call on every rank.  A wait longer than SBLAS_BARRIER_TIMEOUT_MS (environment,
 * default 20000, 0 = no limit) writes 1+peer into *d_status (may be NULL; sticky — the caller must read
 * it before trusting the replica: a timed-out barrier leaves it INCOMPLETE) and returns instead of
 * hanging; with SBLAS_BARRIER_TRAP=1 it also traps, so the stream itself reports the failure.
 * At most 15 ranks. */
SBLAS_API int sblas_peer_barrier(void* stream, int rank, int world, int32_t* d_local_flags,
                                 void* const* d_peer_flags, int32_t epoch, int32_t* d_status);

/* Carry fix-up for rows shared between neighbouring segments (used by the multi-process
 * path too): for each of the n_shared rows, out[row] = beta*out[row] + alpha*sum(carries),
 * carries added in GPU order.  carry holds n_gpu*2*k values ([gpu][head|tail][k]),
 * rows/first/last have n_shared entries: the row index and the first/last contributing GPU. */
SBLAS_API int sblas_carry_fixup_f64(void* stream, int32_t n_shared, const int32_t* h_rows,
                                    const int32_t* h_first_gpu, const int32_t* h_last_gpu,
                                    const double* d_carry, int32_t k, int32_t layout, double* d_out,
                                    int64_t ld, double alpha, double beta);
SBLAS_API int sblas_carry_fixup_f32(void* stream, int32_t n_shared, const int32_t* h_rows,
                                    const int32_t* h_first_gpu, const int32_t* h_last_gpu,
                                    const float* d_carry, int32_t k, int32_t layout, float* d_out,
                                    int64_t ld, float alpha, float beta);

/* ---- synthetic inputs of BASELINE.json's configs, generated on the device --------------
 * (replace matrices/get_matrices.sh:4-5, which needs the network).  All generators are
 * counter-based (splitmix64 keyed by seed and element index) and are restated in numpy in
 * tests/synth_ref.py; the tests check the two bit for bit. */
/* p[i] = lo + (hi-lo)*u01(seed, stream_id, i) */
SBLAS_API int sblas_synth_fill_f64(void* stream, double* d_p, int64_t n, uint64_t seed,
                                   uint64_t stream_id, double lo, double hi);
SBLAS_API int sblas_synth_fill_f32(void* stream, float* d_p, int64_t n, uint64_t seed,
                                   uint64_t stream_id, float lo, float hi);
/* 2-D 5-point Laplacian, nx*ny rows, Dirichlet, columns ascending; nnz = 5*nx*ny-2*nx-2*ny.
 * d_val may be f32 or f64 (val_is_f64). */
SBLAS_API int64_t sblas_synth_laplace2d_nnz(int32_t nx, int32_t ny);
SBLAS_API int sblas_synth_laplace2d(void* stream, int32_t nx, int32_t ny, int32_t* d_rowptr,
                                    int32_t* d_colidx, void* d_val, int val_is_f64);
/* rows x cols, exactly `deg` distinct uniform columns per row (sorted), values in [-1,1). */
SBLAS_API int sblas_synth_uniform(void* stream, int32_t rows, int32_t cols, int32_t deg,
                                  uint64_t seed, int32_t* d_rowptr, int32_t* d_colidx,
                                  void* d_val, int val_is_f64);
/* banded: row i holds the `bw` consecutive columns starting at clamp(i-bw/2, 0, n-bw). */
SBLAS_API int sblas_synth_banded(void* stream, int32_t n, int32_t bw, uint64_t seed,
                                 int32_t* d_rowptr, int32_t* d_colidx, void* d_val,
                                 int val_is_f64);
/* R-MAT (a,b,c,d = .57,.19,.19,.05), 2^scale rows, edge_factor*2^scale edges, duplicates
 * merged (values summed), columns sorted.  Two calls: the first (d_colidx == NULL) returns
 * the merged nnz in *nnz_out and keeps the edges in a device scratch owned by the library;
 * the second fills the three arrays and frees the scratch. */
SBLAS_API int sblas_synth_rmat(void* stream, int32_t scale, int32_t edge_factor, uint64_t seed,
                               int64_t* nnz_out, int32_t* d_rowptr, int32_t* d_colidx,
                               void* d_val, int val_is_f64);

#ifdef __cplusplus
}
#endif
#endif /* SBLAS_B200_H */
