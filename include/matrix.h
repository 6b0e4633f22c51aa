/* include/matrix.h — data model of the SpMV / SpMM path (B200 build, drop-in for matrix.h).
 *
 * Same class names, public field names, constructors and methods as the reference's
 * CsrSparseMatrix (matrix.h:276-453), DenseMatrix (:510-649) and DenseVector (:653-740), and
 * the enums GpuSharePolicy / MajorOrder (:34,:37).  What changes is underneath:
 *   - all memory traffic goes through the C ABI of libsblas_b200.so (utility.h here);
 *   - CsrSparseMatrix::sync2gpu builds, once, a per-GPU execution plan (row statistics and
 *     nnz-balanced merge-path tile tables) that every later SpMV/SpMM call reuses, where the
 *     reference rebuilt cuSPARSE descriptors and work buffers per call (spmv.h:70-101);
 *   - the nnz split keeps the reference's partition tables bit for bit (nnz_gpu,
 *     starting_row_gpu, stoping_row_gpu, local row pointers, matrix.h:360-375) but uses an
 *     integer ceil-div for avg_nnz (the reference's float32 cast drops trailing nnz beyond
 *     2^24) and a binary search instead of the O(height) scan;
 *   - the device copy of a col_major DenseMatrix is stored ROW-major, leading dimension padded
 *     to 16 bytes (dev_layout / ld_gpu below): one nnz of A then selects one contiguous
 *     k*sizeof(T) segment of B.  sync2gpu / sync2cpu convert on the device, so code that
 *     only uses this class sees the reference's col-major semantics.  Set the environment
 *     variable SBLAS_DENSE_DEVICE_LAYOUT=col (or dev_layout before sync2gpu) to keep the
 *     reference's device layout; the ops accept both.
 * COO / CSC containers and CsrToCsc / CscToCsr (matrix.h:40-272, :457-506) keep
 * their names, signatures and field names; the conversions and the COO sort run on GPU 0
 * (sblas_compressed_transpose_on / sblas_coo_sort_on: stable radix sort + gather), bit-identical to
 * the reference's CPU loops.
 * IdxType must be a 32-bit integer type (the only kind the reference's ops are instantiated
 * with, unit_test.cu:81); DataType is float or double. */
#ifndef MATRIX_H
#define MATRIX_H

#include <assert.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "utility.h"

enum GpuSharePolicy { none = 0, replicate = 1, segment = 2 }; /* matrix.h:34 */
enum MajorOrder { row_major = 0, col_major = 1 };             /* matrix.h:37 */

/* ============================ CSR <-> CSC conversion ============================ */
namespace sblas_convert_detail {
/* host arrays in, host arrays out, the work on logical GPU 0 */
template <typename IdxType, typename DataType>
inline void transpose_on_gpu0(IdxType m, IdxType n, IdxType nnz, const IdxType* ptr, const IdxType* idx,
                              const DataType* val, IdxType* out_idx, IdxType* out_ptr, DataType* out_val) {
  static_assert(sizeof(IdxType) == 4, "the B200 build supports 32-bit indices (sblas_b200.h)");
  static_assert(sizeof(DataType) == 4 || sizeof(DataType) == 8, "DataType must be float or double");
  const size_t nz = (size_t)nnz, pb = ((size_t)m + 1) * 4, ob = ((size_t)n + 1) * 4;
  int32_t *d_ptr = NULL, *d_idx = NULL, *d_optr = NULL, *d_oidx = NULL;
  void *d_val = NULL, *d_oval = NULL;
  SAFE_ALOC_GPU_ON(0, d_ptr, pb);
  SAFE_ALOC_GPU_ON(0, d_optr, ob);
  SAFE_ALOC_GPU_ON(0, d_idx, std::max<size_t>(nz * 4, 16));
  SAFE_ALOC_GPU_ON(0, d_oidx, std::max<size_t>(nz * 4, 16));
  SAFE_ALOC_GPU_ON(0, d_val, std::max<size_t>(nz * sizeof(DataType), 16));
  SAFE_ALOC_GPU_ON(0, d_oval, std::max<size_t>(nz * sizeof(DataType), 16));
  SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_ptr, ptr, pb));
  if (nz) {
    SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_idx, idx, nz * 4));
    SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_val, val, nz * sizeof(DataType)));
  }
  SBLAS_SAFE_CALL(sblas_compressed_transpose_on(0, sizeof(DataType) == 8, (int32_t)m, (int32_t)n, (int64_t)nnz, d_ptr,
                                                d_idx, d_val, d_optr, d_oidx, d_oval));
  SBLAS_SAFE_CALL(sblas_copy_d2h(0, out_ptr, d_optr, ob));
  if (nz) {
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, out_idx, d_oidx, nz * 4));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, out_val, d_oval, nz * sizeof(DataType)));
  }
  SAFE_FREE_GPU_ON(0, d_ptr);
  SAFE_FREE_GPU_ON(0, d_optr);
  SAFE_FREE_GPU_ON(0, d_idx);
  SAFE_FREE_GPU_ON(0, d_oidx);
  SAFE_FREE_GPU_ON(0, d_val);
  SAFE_FREE_GPU_ON(0, d_oval);
}
}  // namespace sblas_convert_detail

/* matrix.h:40-64 — same signature (host arrays); inside a column the entries keep CSR order */
template <typename IdxType, typename DataType>
void CsrToCsc(const IdxType m, const IdxType n, const IdxType nnz, const IdxType* csrRowPtr, const IdxType* csrColIdx,
              const DataType* csrVal, IdxType* cscRowIdx, IdxType* cscColPtr, DataType* cscVal) {
  sblas_convert_detail::transpose_on_gpu0<IdxType, DataType>(m, n, nnz, csrRowPtr, csrColIdx, csrVal, cscRowIdx,
                                                            cscColPtr, cscVal);
}
/* matrix.h:67-91 — note the reference's argument order: (n, m, ...) */
template <typename IdxType, typename DataType>
void CscToCsr(const IdxType n, const IdxType m, const IdxType nnz, const IdxType* cscColPtr, const IdxType* cscRowIdx,
              const DataType* cscVal, IdxType* csrColIdx, IdxType* csrRowPtr, DataType* csrVal) {
  sblas_convert_detail::transpose_on_gpu0<IdxType, DataType>(n, m, nnz, cscColPtr, cscRowIdx, cscVal, csrColIdx,
                                                            csrRowPtr, csrVal);
}

/* ================================== COO matrix ================================== */
/* matrix.h:117-272.  Loaded from a Matrix Market file as the file holds it (no symmetric expansion,
 * matrix.h:165-173), sorted by (row, col) on the device, then replicated or cut into
 * floor(nnz / n_gpu)-sized pieces exactly like matrix.h:196-216 (trailing nnz % n_gpu entries stay
 * on the host only, as in the reference). */
template <typename IdxType, typename DataType>
class CooSparseMatrix {
  static_assert(sizeof(IdxType) == 4, "the B200 build supports 32-bit indices (sblas_b200.h)");

 public:
  CooSparseMatrix() { clear_(); }
  explicit CooSparseMatrix(const char* filename, unsigned _n_gpu = 0, enum GpuSharePolicy _policy = none) {
    clear_();
    std::cout << "Loading input matrix from '" << filename << "'." << std::endl;
    int32_t m = 0, n = 0, nz = 0;
    SBLAS_SAFE_CALL(sblas_mtx_coo_info(filename, &m, &n, &nz));
    height = (IdxType)m;
    width = (IdxType)n;
    nnz = (IdxType)nz;
    std::cout << "Height: " << height << " Width: " << width << " nnz: " << nnz << std::endl;
    SAFE_ALOC_HOST(cooRowIdx, std::max<size_t>(get_nnz_idx_size(), 16));
    SAFE_ALOC_HOST(cooColIdx, std::max<size_t>(get_nnz_idx_size(), 16));
    SAFE_ALOC_HOST(cooVal, std::max<size_t>(get_nnz_val_size(), 16));
    double* tmp = (double*)malloc(sizeof(double) * (size_t)std::max(nz, 1));
    SBLAS_SAFE_CALL(sblas_mtx_coo_data(filename, (int32_t*)cooRowIdx, (int32_t*)cooColIdx, tmp));
    for (int32_t j = 0; j < nz; j++) cooVal[j] = (DataType)tmp[j];
    free(tmp);
    sortByRow();
    n_gpu = _n_gpu;
    policy = _policy;
    if (n_gpu != 0 && policy != none) {
      SAFE_ALOC_HOST(cooRowIdx_gpu, n_gpu * sizeof(IdxType*));
      SAFE_ALOC_HOST(cooColIdx_gpu, n_gpu * sizeof(IdxType*));
      SAFE_ALOC_HOST(cooVal_gpu, n_gpu * sizeof(DataType*));
      SAFE_ALOC_HOST(nnz_gpu, n_gpu * sizeof(IdxType));
      const IdxType avg_nnz = nnz / (IdxType)n_gpu; /* matrix.h:198: ceil() of an integer quotient */
      for (unsigned i = 0; i < n_gpu; i++) {
        const IdxType first = policy == segment ? (IdxType)i * avg_nnz : 0;
        nnz_gpu[i] = policy == segment ? std::min((IdxType)(i + 1) * avg_nnz, nnz) - first : nnz;
        cooRowIdx_gpu[i] = cooColIdx_gpu[i] = NULL;
        cooVal_gpu[i] = NULL;
        SAFE_ALOC_GPU_ON(i, cooRowIdx_gpu[i], std::max<size_t>(get_gpu_nnz_idx_size(i), 16));
        SAFE_ALOC_GPU_ON(i, cooColIdx_gpu[i], std::max<size_t>(get_gpu_nnz_idx_size(i), 16));
        SAFE_ALOC_GPU_ON(i, cooVal_gpu[i], std::max<size_t>(get_gpu_nnz_val_size(i), 16));
        if (nnz_gpu[i] > 0) {
          SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, cooRowIdx_gpu[i], cooRowIdx + first, get_gpu_nnz_idx_size(i)));
          SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, cooColIdx_gpu[i], cooColIdx + first, get_gpu_nnz_idx_size(i)));
          SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, cooVal_gpu[i], cooVal + first, get_gpu_nnz_val_size(i)));
        }
      }
    }
  }
  ~CooSparseMatrix() {
    SAFE_FREE_HOST(cooRowIdx);
    SAFE_FREE_HOST(cooColIdx);
    SAFE_FREE_HOST(cooVal);
    if (cooRowIdx_gpu != NULL) {
      for (unsigned i = 0; i < n_gpu; i++) {
        SAFE_FREE_GPU_ON(i, cooRowIdx_gpu[i]);
        SAFE_FREE_GPU_ON(i, cooColIdx_gpu[i]);
        SAFE_FREE_GPU_ON(i, cooVal_gpu[i]);
      }
      SAFE_FREE_HOST(cooRowIdx_gpu);
      SAFE_FREE_HOST(cooColIdx_gpu);
      SAFE_FREE_HOST(cooVal_gpu);
    }
    SAFE_FREE_HOST(nnz_gpu);
  }
  /* matrix.h:221-237 — by (row, col); equal pairs keep file order */
  void sortByRow() {
    if (nnz <= 0) return;
    int32_t *d_r = NULL, *d_c = NULL;
    void* d_v = NULL;
    SAFE_ALOC_GPU_ON(0, d_r, get_nnz_idx_size());
    SAFE_ALOC_GPU_ON(0, d_c, get_nnz_idx_size());
    SAFE_ALOC_GPU_ON(0, d_v, get_nnz_val_size());
    SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_r, cooRowIdx, get_nnz_idx_size()));
    SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_c, cooColIdx, get_nnz_idx_size()));
    SBLAS_SAFE_CALL(sblas_copy_h2d(0, d_v, cooVal, get_nnz_val_size()));
    SBLAS_SAFE_CALL(sblas_coo_sort_on(0, sizeof(DataType) == 8, (int64_t)nnz, (int32_t)height, d_r, d_c, d_v));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, cooRowIdx, d_r, get_nnz_idx_size()));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, cooColIdx, d_c, get_nnz_idx_size()));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, cooVal, d_v, get_nnz_val_size()));
    SAFE_FREE_GPU_ON(0, d_r);
    SAFE_FREE_GPU_ON(0, d_c);
    SAFE_FREE_GPU_ON(0, d_v);
  }
  size_t get_gpu_nnz_idx_size(unsigned i_gpu) {
    assert(i_gpu < n_gpu);
    return nnz_gpu != NULL ? (size_t)nnz_gpu[i_gpu] * sizeof(IdxType) : 0;
  }
  size_t get_gpu_nnz_val_size(unsigned i_gpu) {
    assert(i_gpu < n_gpu);
    return nnz_gpu != NULL ? (size_t)nnz_gpu[i_gpu] * sizeof(DataType) : 0;
  }
  size_t get_nnz_idx_size() { return (size_t)nnz * sizeof(IdxType); }
  size_t get_nnz_val_size() { return (size_t)nnz * sizeof(DataType); }

 public: /* matrix.h:255-271 */
  IdxType* cooRowIdx;
  IdxType* cooColIdx;
  DataType* cooVal;
  IdxType** cooRowIdx_gpu;
  IdxType** cooColIdx_gpu;
  DataType** cooVal_gpu;
  IdxType* nnz_gpu;
  IdxType nnz;
  IdxType height;
  IdxType width;
  unsigned n_gpu;
  enum GpuSharePolicy policy;

 private:
  void clear_() {
    cooRowIdx = cooColIdx = NULL;
    cooVal = NULL;
    cooRowIdx_gpu = cooColIdx_gpu = NULL;
    cooVal_gpu = NULL;
    nnz_gpu = NULL;
    nnz = height = width = 0;
    n_gpu = 0;
    policy = none;
  }
};

/* ================================== CSR matrix ================================== */
template <typename IdxType, typename DataType>
class CsrSparseMatrix {
  static_assert(sizeof(IdxType) == 4, "the B200 build supports 32-bit indices (sblas_b200.h)");

 public:
  CsrSparseMatrix() { clear_(); }

  /* matrix.h:300-330 — load a Matrix Market file (pattern/real/integer, general/symmetric;
   * entries of a row stay in file order, exactly like mmio_highlevel.h:130-281). */
  explicit CsrSparseMatrix(const char* filename) {
    clear_();
    int32_t m = 0, n = 0, nnzA = 0, sym = 0;
    SBLAS_SAFE_CALL(sblas_mtx_info(filename, &m, &n, &nnzA, &sym));
    height = (IdxType)m;
    width = (IdxType)n;
    nnz = (IdxType)nnzA;
    SAFE_ALOC_HOST(csrRowPtr, get_row_ptr_size());
    SAFE_ALOC_HOST(csrColIdx, get_col_idx_size());
    SAFE_ALOC_HOST(csrVal, get_val_size());
    double* tmp = (double*)malloc(sizeof(double) * (size_t)std::max(nnzA, 1));
    SBLAS_SAFE_CALL(sblas_mtx_data(filename, (int32_t*)csrRowPtr, (int32_t*)csrColIdx, tmp));
    for (int32_t j = 0; j < nnzA; j++) csrVal[j] = (DataType)tmp[j];
    free(tmp);
    printf("input matrix A: ( %i, %i ) nnz = %i\n", m, n, nnzA);
  }

  /* Not in the reference: adopt host arrays (copied), for synthetic matrices. */
  CsrSparseMatrix(IdxType h, IdxType w, const IdxType* rowptr, const IdxType* colidx, const DataType* val) {
    clear_();
    height = h;
    width = w;
    nnz = rowptr[h];
    SAFE_ALOC_HOST(csrRowPtr, get_row_ptr_size());
    SAFE_ALOC_HOST(csrColIdx, get_col_idx_size());
    SAFE_ALOC_HOST(csrVal, get_val_size());
    memcpy(csrRowPtr, rowptr, get_row_ptr_size());
    memcpy(csrColIdx, colidx, get_col_idx_size());
    memcpy(csrVal, val, get_val_size());
  }

  ~CsrSparseMatrix() {
    release_gpu_();
    SAFE_FREE_HOST(csrRowPtr);
    SAFE_FREE_HOST(csrColIdx);
    SAFE_FREE_HOST(csrVal);
  }
  CsrSparseMatrix(const CsrSparseMatrix&) = delete;
  CsrSparseMatrix& operator=(const CsrSparseMatrix&) = delete;

  /* matrix.h:331-397 */
  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) {
    sblas_range range__("sblas sync2gpu (partition + upload)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    release_gpu_();
    n_gpu = _n_gpu;
    policy = _policy;
    SAFE_ALOC_HOST(csrRowPtr_gpu, n_gpu * sizeof(IdxType*));
    SAFE_ALOC_HOST(csrColIdx_gpu, n_gpu * sizeof(IdxType*));
    SAFE_ALOC_HOST(csrVal_gpu, n_gpu * sizeof(DataType*));
    SAFE_ALOC_HOST(plan_gpu, n_gpu * sizeof(sblasPlan_t));
    for (unsigned i = 0; i < n_gpu; i++) {
      csrRowPtr_gpu[i] = NULL;
      csrColIdx_gpu[i] = NULL;
      csrVal_gpu[i] = NULL;
      plan_gpu[i] = NULL;
    }
    if (policy == replicate) {
      for (unsigned i = 0; i < n_gpu; i++)
        upload_(i, csrRowPtr, (size_t)height + 1, csrColIdx, csrVal, (size_t)nnz);
    } else {
      SAFE_ALOC_HOST(nnz_gpu, n_gpu * sizeof(IdxType));
      SAFE_ALOC_HOST(starting_row_gpu, n_gpu * sizeof(IdxType));
      SAFE_ALOC_HOST(stoping_row_gpu, n_gpu * sizeof(IdxType));
      int32_t avg_nnz = 0;
      SBLAS_SAFE_CALL(sblas_csr_partition_nnz_i32((const int32_t*)csrRowPtr, (int32_t)height, (int32_t)nnz, n_gpu,
                                                  &avg_nnz, (int32_t*)nnz_gpu, (int32_t*)starting_row_gpu,
                                                  (int32_t*)stoping_row_gpu));
      for (unsigned i = 0; i < n_gpu; i++) {
        const bool has_rows = nnz_gpu[i] > 0 && starting_row_gpu[i] >= 0 && stoping_row_gpu[i] >= starting_row_gpu[i];
        if (has_rows) {
          const size_t np = get_gpu_row_ptr_num(i);
          IdxType* fixedRowPtr = (IdxType*)malloc(np * sizeof(IdxType));
          SBLAS_SAFE_CALL(sblas_csr_segment_rowptr_i32((const int32_t*)csrRowPtr, avg_nnz, i,
                                                       (int32_t)starting_row_gpu[i], (int32_t)stoping_row_gpu[i],
                                                       (int32_t)nnz_gpu[i], (int32_t*)fixedRowPtr));
          const size_t off = (size_t)i * (size_t)avg_nnz;
          upload_(i, fixedRowPtr, np, csrColIdx + off, csrVal + off, (size_t)nnz_gpu[i]);
          free(fixedRowPtr);
        }
        printf("gpu-%d,start-row:%d,stop-row:%d,num-rows:%ld,num-nnz:%d\n", i, (int)starting_row_gpu[i],
               (int)stoping_row_gpu[i], (long)get_gpu_row_ptr_num(i), (int)nnz_gpu[i]); /* matrix.h:388 */
      }
    }
  }

  /* sizes, as in matrix.h:398-431 */
  size_t get_gpu_row_ptr_num(unsigned i_gpu) const {
    assert(i_gpu < n_gpu);
    return nnz_gpu != NULL ? (size_t)(stoping_row_gpu[i_gpu] - starting_row_gpu[i_gpu] + 2) : 0;
  }
  size_t get_gpu_row_ptr_size(unsigned i_gpu) const { return get_gpu_row_ptr_num(i_gpu) * sizeof(IdxType); }
  size_t get_gpu_col_idx_num(unsigned i_gpu) const {
    assert(i_gpu < n_gpu);
    return nnz_gpu != NULL ? (size_t)nnz_gpu[i_gpu] : 0;
  }
  size_t get_gpu_col_idx_size(unsigned i_gpu) const { return get_gpu_col_idx_num(i_gpu) * sizeof(IdxType); }
  size_t get_gpu_nnz_val_num(unsigned i_gpu) const { return get_gpu_col_idx_num(i_gpu); }
  size_t get_gpu_nnz_val_size(unsigned i_gpu) const { return get_gpu_nnz_val_num(i_gpu) * sizeof(DataType); }
  size_t get_row_ptr_size() const { return ((size_t)height + 1) * sizeof(IdxType); }
  size_t get_col_idx_size() const { return (size_t)nnz * sizeof(IdxType); }
  size_t get_val_size() const { return (size_t)nnz * sizeof(DataType); }

 public: /* the reference's public fields, matrix.h:433-452 */
  IdxType* csrRowPtr;
  IdxType* csrColIdx;
  DataType* csrVal;
  IdxType** csrRowPtr_gpu;
  IdxType** csrColIdx_gpu;
  DataType** csrVal_gpu;
  IdxType* nnz_gpu;          /* nnz of each GPU's segment */
  IdxType* starting_row_gpu; /* first / last row touched by the segment: a row may be shared */
  IdxType* stoping_row_gpu;  /* by neighbouring GPUs (matrix.h:443-445) */
  IdxType nnz;
  IdxType height;
  IdxType width;
  unsigned n_gpu;
  enum GpuSharePolicy policy;
  /* B200 build: per-GPU execution plans, built by sync2gpu (NULL where a GPU owns nothing) */
  sblasPlan_t* plan_gpu;

 private:
  void clear_() {
    csrRowPtr = csrColIdx = NULL;
    csrVal = NULL;
    csrRowPtr_gpu = csrColIdx_gpu = NULL;
    csrVal_gpu = NULL;
    nnz_gpu = starting_row_gpu = stoping_row_gpu = NULL;
    plan_gpu = NULL;
    nnz = height = width = 0;
    n_gpu = 0;
    policy = none;
  }
  void upload_(unsigned i, const IdxType* rp, size_t n_rp, const IdxType* ci, const DataType* v, size_t n_nz) {
    SAFE_ALOC_GPU_ON(i, csrRowPtr_gpu[i], n_rp * sizeof(IdxType));
    SAFE_ALOC_GPU_ON(i, csrColIdx_gpu[i], n_nz * sizeof(IdxType));
    SAFE_ALOC_GPU_ON(i, csrVal_gpu[i], n_nz * sizeof(DataType));
    SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, csrRowPtr_gpu[i], rp, n_rp * sizeof(IdxType)));
    SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, csrColIdx_gpu[i], ci, n_nz * sizeof(IdxType)));
    SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, csrVal_gpu[i], v, n_nz * sizeof(DataType)));
    SBLAS_SAFE_CALL(sblas_plan_create_on((int)i, (int32_t)(n_rp - 1), (int64_t)n_nz,
                                         (const int32_t*)csrRowPtr_gpu[i], &plan_gpu[i]));
    /* SpMM row-block mirror of this GPU's rows (kept only for structured inputs, sblas_b200.h) */
    SBLAS_SAFE_CALL(sblas_plan_build_rowblocks_on((int)i, sizeof(DataType) == 8, plan_gpu[i],
                                                  (const int32_t*)csrColIdx_gpu[i], csrVal_gpu[i]));
    /* SpMV: tile table + column-locality statistic now (no SpMV call then synchronises), and the column-panel
       mirror of this GPU's rows when x will not stay in L2 (gather-heavy inputs only, sblas_b200.h) */
    SBLAS_SAFE_CALL(sblas_plan_prepare_spmv_on((int)i, sizeof(DataType) == 8, plan_gpu[i],
                                               (const int32_t*)csrColIdx_gpu[i], csrVal_gpu[i], (int32_t)width));
  }
  void release_gpu_() {
    if (plan_gpu != NULL) {
      for (unsigned i = 0; i < n_gpu; i++)
        if (plan_gpu[i] != NULL) sblas_plan_destroy(plan_gpu[i]);
      SAFE_FREE_HOST(plan_gpu);
    }
    SAFE_FREE_MULTI_GPU(csrRowPtr_gpu, n_gpu);
    SAFE_FREE_MULTI_GPU(csrColIdx_gpu, n_gpu);
    SAFE_FREE_MULTI_GPU(csrVal_gpu, n_gpu);
    SAFE_FREE_HOST(nnz_gpu);
    SAFE_FREE_HOST(starting_row_gpu);
    SAFE_FREE_HOST(stoping_row_gpu);
  }
};

/* ================================== CSC matrix ================================== */
/* matrix.h:457-506 — built from a CSR matrix by CsrToCsc (on the device, see above) */
template <typename IdxType, typename DataType>
class CscSparseMatrix {
 public:
  CscSparseMatrix() : cscRowIdx(NULL), cscColPtr(NULL), cscVal(NULL), nnz(0), height(0), width(0) {}
  explicit CscSparseMatrix(const CsrSparseMatrix<IdxType, DataType>* csr)
      : cscRowIdx(NULL), cscColPtr(NULL), cscVal(NULL) {
    height = csr->height;
    width = csr->width;
    nnz = csr->nnz;
    std::cout << "Building csc matrix from a csr matrix." << std::endl;
    std::cout << "Height: " << height << " Width: " << width << " nnz: " << nnz << std::endl;
    SAFE_ALOC_HOST(cscColPtr, get_col_ptr_size());
    SAFE_ALOC_HOST(cscRowIdx, std::max<size_t>(get_row_idx_size(), 16));
    SAFE_ALOC_HOST(cscVal, std::max<size_t>(get_val_size(), 16));
    CsrToCsc<IdxType, DataType>(height, width, nnz, csr->csrRowPtr, csr->csrColIdx, csr->csrVal, cscRowIdx, cscColPtr,
                                cscVal);
  }
  ~CscSparseMatrix() {
    SAFE_FREE_HOST(cscColPtr);
    SAFE_FREE_HOST(cscRowIdx);
    SAFE_FREE_HOST(cscVal);
  }
  size_t get_col_ptr_size() { return ((size_t)width + 1) * sizeof(IdxType); }
  size_t get_row_idx_size() { return (size_t)nnz * sizeof(IdxType); }
  size_t get_val_size() { return (size_t)nnz * sizeof(DataType); }

 public: /* matrix.h:499-505 */
  IdxType* cscRowIdx;
  IdxType* cscColPtr;
  DataType* cscVal;
  IdxType nnz;
  IdxType height;
  IdxType width;
};

/* device layout a col_major DenseMatrix gets on sync2gpu unless overridden per object */
inline int32_t sblas_default_dense_device_layout() {
  const char* e = getenv("SBLAS_DENSE_DEVICE_LAYOUT");
  return (e != NULL && (e[0] == 'c' || e[0] == 'C')) ? SBLAS_COL_MAJOR : SBLAS_ROW_MAJOR;
}

/* ================================== dense matrix ================================== */
template <typename IdxType, typename DataType>
class DenseMatrix {
 public:
  DenseMatrix() : height(0), width(0), order(row_major) { init_(); }

  /* matrix.h:519-529 — values rand()/RAND_MAX in storage order after srand(RAND_INIT_SEED) */
  DenseMatrix(IdxType _height, IdxType _width, enum MajorOrder _order)
      : height(_height), width(_width), order(_order) {
    init_();
    SAFE_ALOC_HOST(val, get_mtx_size());
    srand(RAND_INIT_SEED);
    const size_t n = get_mtx_num();
    for (size_t i = 0; i < n; i++) val[i] = (DataType)rand0to1();
  }
  /* matrix.h:530-539 — constant fill */
  DenseMatrix(IdxType _height, IdxType _width, DataType _val, enum MajorOrder _order)
      : height(_height), width(_width), order(_order) {
    init_();
    SAFE_ALOC_HOST(val, get_mtx_size());
    srand(RAND_INIT_SEED);
    std::fill(val, val + get_mtx_num(), _val);
  }
  ~DenseMatrix() {
    release_gpu_();
    SAFE_FREE_HOST(val);
  }
  DenseMatrix(const DenseMatrix&) = delete;
  DenseMatrix& operator=(const DenseMatrix&) = delete;

  /* matrix.h:541-570.  replicate: every GPU gets the whole matrix; segment: GPU i gets
   * dim_gpu[i] lines of the leading dimension (columns for col_major), contiguous on the host. */
  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) {
    sblas_range range__("sblas sync2gpu (partition + upload)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    release_gpu_();
    n_gpu = _n_gpu;
    policy = _policy;
    /* only col_major host matrices are re-laid out; a row_major one already has the layout */
    if (order == row_major) dev_layout = SBLAS_COL_MAJOR; /* "keep host layout" in upload terms */
    else if (dev_layout < 0) dev_layout = sblas_default_dense_device_layout();
    const IdxType first_order = (order == row_major) ? height : width;
    const IdxType second_order = (order == row_major) ? width : height;
    SAFE_ALOC_HOST(val_gpu, n_gpu * sizeof(DataType*));
    SAFE_ALOC_HOST(ld_gpu, n_gpu * sizeof(int64_t));
    int32_t avg = (int32_t)first_order;
    if (policy == segment) {
      SAFE_ALOC_HOST(dim_gpu, n_gpu * sizeof(IdxType));
      SBLAS_SAFE_CALL(sblas_dense_split_i32((int32_t)first_order, n_gpu, &avg, (int32_t*)dim_gpu));
    }
    avg_val = (IdxType)avg;
    for (unsigned i = 0; i < n_gpu; i++) {
      const int64_t dim = (policy == segment) ? (int64_t)dim_gpu[i] : (int64_t)first_order;
      const size_t off = (policy == segment) ? (size_t)i * (size_t)avg * (size_t)second_order : 0;
      val_gpu[i] = NULL;
      ld_gpu[i] = second_order;
      if (dim > 0) /* trailing GPUs may own nothing */
        SBLAS_SAFE_CALL(sblas_abi<DataType>::upload((int)i, val + off, (int64_t)second_order, dim, dev_layout,
                                                    &val_gpu[i], &ld_gpu[i]));
    }
  }

  /* matrix.h:595-612 — copy GPU i's slice (segment) or the whole matrix (replicate) back */
  void sync2cpu(unsigned i_gpu) {
    assert(val_gpu != NULL);
    assert(i_gpu < n_gpu);
    const IdxType first_order = (order == row_major) ? height : width;
    const IdxType second_order = (order == row_major) ? width : height;
    const int64_t dim = (policy == segment) ? (int64_t)dim_gpu[i_gpu] : (int64_t)first_order;
    const size_t off = (policy == segment) ? (size_t)i_gpu * (size_t)avg_val * (size_t)second_order : 0;
    if (dim > 0)
      SBLAS_SAFE_CALL(sblas_abi<DataType>::download((int)i_gpu, val_gpu[i_gpu], ld_gpu[i_gpu], (int64_t)second_order,
                                                    dim, dev_layout, val + off));
  }

  /* B200 build: sync2cpu(i) for i < n, with the downloads of all GPUs in flight together (each GPU has its own host
   * link); what sblas_spmm_csr_v1 uses for its in-op refresh of the host copy (spmm.h:151) */
  void sync2cpu_all(unsigned n) {
    assert(val_gpu != NULL);
    assert(n <= n_gpu);
    if (policy != segment) { if (n > 0) sync2cpu(0); return; }
    const IdxType second_order = (order == row_major) ? width : height;
    const DataType* d[SBLAS_MAX_GPU];
    DataType* h[SBLAS_MAX_GPU];
    int64_t dims[SBLAS_MAX_GPU];
    for (unsigned i = 0; i < n; i++) {
      d[i] = val_gpu[i];
      dims[i] = (int64_t)dim_gpu[i];
      h[i] = val + (size_t)i * (size_t)avg_val * (size_t)second_order;
    }
    SBLAS_SAFE_CALL(sblas_abi<DataType>::download_multi(n, d, ld_gpu, (int64_t)second_order, dims, dev_layout, h));
  }

  /* matrix.h:577-593 — host transpose into a new matrix of the other major order */
  DenseMatrix* transpose() {
    assert(n_gpu == 0);
    assert(policy == none);
    DenseMatrix* t = new DenseMatrix(height, width, DataType(0), order == row_major ? col_major : row_major);
    for (IdxType i = 0; i < height; i++)
      for (IdxType j = 0; j < width; j++) {
        const size_t rm = (size_t)i * width + j, cm = (size_t)j * height + i;
        if (order == row_major) t->val[cm] = val[rm]; else t->val[rm] = val[cm];
      }
    return t;
  }

  /* matrix.h:613-625 — this = this*beta + dm*alpha on every GPU (same policy/layout required) */
  void plusDenseMatrixGPU(DenseMatrix const& dm, DataType alpha, DataType beta) {
    if (n_gpu == 0 || policy == none) return;
    assert(dm.n_gpu >= n_gpu && dm.dev_layout == dev_layout);
    for (unsigned i = 0; i < n_gpu; i++) {
      const int64_t n = device_elems_(i);
      if (n > 0) SBLAS_SAFE_CALL(sblas_axpby_on((int)i, sizeof(DataType) == 8, val_gpu[i], dm.val_gpu[i],
                                                (double)alpha, (double)beta, n));
    }
    sblas_sync_gpus(n_gpu);
  }

  size_t get_dim_gpu_num(unsigned i_gpu) const {
    assert(i_gpu < n_gpu);
    return (size_t)dim_gpu[i_gpu];
  }
  size_t get_dim_gpu_size(unsigned i_gpu) const { return get_dim_gpu_num(i_gpu) * sizeof(DataType); }
  size_t get_row_size() const { return (size_t)width * sizeof(DataType); }
  size_t get_col_size() const { return (size_t)height * sizeof(DataType); }
  size_t get_mtx_num() const { return (size_t)width * (size_t)height; }
  size_t get_mtx_size() const { return get_mtx_num() * sizeof(DataType); }

 public: /* the reference's public fields, matrix.h:638-648 */
  IdxType height;
  IdxType width;
  DataType* val;
  DataType** val_gpu;
  IdxType* dim_gpu; /* lines of the leading dimension per GPU (segment policy) */
  unsigned n_gpu;
  enum GpuSharePolicy policy;
  enum MajorOrder order;
  /* B200 build: layout of the device copies (SBLAS_ROW_MAJOR / SBLAS_COL_MAJOR; -1 = take the
   * default at sync2gpu) and their leading dimensions */
  int32_t dev_layout;
  int64_t* ld_gpu;
  IdxType avg_val;

 private:
  void init_() {
    val = NULL;
    val_gpu = NULL;
    dim_gpu = NULL;
    ld_gpu = NULL;
    n_gpu = 0;
    policy = none;
    dev_layout = -1;
    avg_val = 0;
  }
  int64_t device_elems_(unsigned i) const {
    const IdxType first_order = (order == row_major) ? height : width;
    const IdxType second_order = (order == row_major) ? width : height;
    const int64_t dim = (policy == segment) ? (int64_t)dim_gpu[i] : (int64_t)first_order;
    if (dim <= 0) return 0;
    return (order == col_major && dev_layout == SBLAS_ROW_MAJOR) ? (int64_t)second_order * ld_gpu[i]
                                                                  : (int64_t)second_order * dim;
  }
  void release_gpu_() {
    SAFE_FREE_MULTI_GPU(val_gpu, n_gpu);
    SAFE_FREE_HOST(dim_gpu);
    SAFE_FREE_HOST(ld_gpu);
    n_gpu = 0;
    policy = none;
  }
};

/* ================================== dense vector ================================== */
template <typename IdxType, typename DataType>
class DenseVector {
 public:
  DenseVector() : length(0), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {}
  /* matrix.h:659-665 — random fill, seed RAND_INIT_SEED */
  explicit DenseVector(IdxType _length) : length(_length), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {
    SAFE_ALOC_HOST(val, get_vec_size());
    srand(RAND_INIT_SEED);
    for (size_t i = 0; i < get_vec_length(); i++) val[i] = (DataType)rand0to1();
  }
  /* matrix.h:666-673 — constant fill */
  DenseVector(IdxType _length, DataType _val) : length(_length), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {
    SAFE_ALOC_HOST(val, get_vec_size());
    srand(RAND_INIT_SEED);
    std::fill(val, val + get_vec_length(), _val);
  }
  /* matrix.h:674-687 — deep copy, device replicas included */
  DenseVector(const DenseVector& dv) : length(dv.length), val(NULL), val_gpu(NULL), n_gpu(dv.n_gpu), policy(dv.policy) {
    SAFE_ALOC_HOST(val, get_vec_size());
    memcpy(val, dv.val, get_vec_size());
    if (n_gpu != 0 && policy != none) {
      SAFE_ALOC_HOST(val_gpu, n_gpu * sizeof(DataType*));
      for (unsigned i = 0; i < n_gpu; i++) {
        val_gpu[i] = NULL;
        SAFE_ALOC_GPU_ON(i, val_gpu[i], get_vec_size());
        SBLAS_SAFE_CALL(sblas_copy_d2d((int)i, val_gpu[i], dv.val_gpu[i], get_vec_size()));
      }
    }
  }
  DenseVector& operator=(const DenseVector&) = delete;
  ~DenseVector() {
    SAFE_FREE_MULTI_GPU(val_gpu, n_gpu);
    SAFE_FREE_HOST(val);
  }

  /* matrix.h:688-704 — vectors are always replicated */
  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) {
    sblas_range range__("sblas sync2gpu (partition + upload)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    assert(_policy != segment);
    SAFE_FREE_MULTI_GPU(val_gpu, n_gpu);
    n_gpu = _n_gpu;
    policy = _policy;
    SAFE_ALOC_HOST(val_gpu, n_gpu * sizeof(DataType*));
    for (unsigned i = 0; i < n_gpu; i++) {
      val_gpu[i] = NULL;
      SAFE_ALOC_GPU_ON(i, val_gpu[i], get_vec_size());
      SBLAS_SAFE_CALL(sblas_copy_h2d((int)i, val_gpu[i], val, get_vec_size()));
    }
  }
  /* matrix.h:705-713 — every GPU holds the same result; take it from any of them */
  void sync2cpu(unsigned i_gpu) {
    assert(i_gpu < n_gpu);
    assert(val_gpu != NULL);
    SBLAS_SAFE_CALL(sblas_copy_d2h((int)i_gpu, val, val_gpu[i_gpu], get_vec_size()));
  }
  /* matrix.h:714-726 — this = this*beta + dv*alpha on every GPU */
  void plusDenseVectorGPU(DenseVector const& dv, DataType alpha, DataType beta) {
    if (n_gpu == 0 || policy == none) return;
    for (unsigned i = 0; i < n_gpu; i++)
      SBLAS_SAFE_CALL(sblas_axpby_on((int)i, sizeof(DataType) == 8, val_gpu[i], dv.val_gpu[i], (double)alpha,
                                     (double)beta, (int64_t)get_vec_length()));
    sblas_sync_gpus(n_gpu);
  }
  size_t get_vec_size() const { return (size_t)length * sizeof(DataType); }
  size_t get_vec_length() const { return (size_t)length; }

 public: /* matrix.h:735-739 */
  IdxType length;
  DataType* val;
  DataType** val_gpu;
  unsigned n_gpu;
  enum GpuSharePolicy policy;
};

#endif /* MATRIX_H */
