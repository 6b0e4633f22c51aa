/* include/matrix_um.h — managed-memory data model of the SpMV / SpMM path (B200 build,
 * drop-in for the reference's matrix_um.h; do not include together with matrix.h: the class
 * names are the same, as in the reference).
 *
 * Semantics kept from the reference (matrix_um.h:384-672, :721-927, :931-1075):
 *   - every array is one cudaMallocManaged allocation; "sync2gpu" copies nothing for GPU 0 —
 *     replicate: replica 0 IS the base array, replicas 1.. are managed copies made with memcpy
 *     (:470-483, :771-778, :980-993); segment: col_idx / val / dense slices are POINTERS INTO the
 *     base array (:517-518, :788), only the per-GPU row pointers are new arrays (:506-513);
 *   - sync2cpu moves nothing: optional prefetch to the CPU + device synchronise (:870-890,
 *     :1031-1045), because the result already lives in the base array (replica 0 / the slices);
 *   - applyGpuTuning / removeGpuTuning / applyCpuTuning set cudaMemAdvise hints
 *     (ReadMostly for inputs, PreferredLocation for outputs) and prefetch the per-GPU ranges
 *     (:529-614, :793-838, :1001-1029).
 * B200 build specifics: dense operands stay col-major on the device (they alias the host
 * arrays), so the ops run the native col-major SpMM kernel; segment slices of col_idx / val
 * start at i*avg_nnz and are generally not 16-byte aligned, which the kernels detect (they take
 * the element-wise load path instead of the TMA bulk path).  NVTX ranges as in the reference
 * (:435-445 etc.): sync2gpu, the tuning hooks and the ops push / pop named ranges (sblas_range_push). */
#ifndef MATRIX_UM_H
#define MATRIX_UM_H
#ifdef MATRIX_H
#error "matrix.h and matrix_um.h define the same class names: include only one of them"
#endif

#include <assert.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "utility.h"

enum GpuSharePolicy { none = 0, replicate = 1, segment = 2 };
enum MajorOrder { row_major = 0, col_major = 1 };

namespace sblas_um_detail {
/* advise + prefetch one managed range towards logical GPU `gpu` (applyGpuTuning), or drop the
 * advice and (optionally) pull it back to the CPU (removeGpuTuning) */
inline void tune_range(void* p, size_t bytes, bool readOnly, int gpu) {
  if (p == NULL || bytes == 0) return;
  SBLAS_SAFE_CALL(sblas_mem_advise(p, bytes, readOnly ? SBLAS_ADVISE_SET_READ_MOSTLY : SBLAS_ADVISE_SET_PREFERRED_LOCATION, gpu));
  SBLAS_SAFE_CALL(sblas_mem_prefetch(p, bytes, gpu));
}
inline void untune_range(void* p, size_t bytes, bool readOnly, int gpu, bool prefetch_cpu) {
  if (p == NULL || bytes == 0) return;
  SBLAS_SAFE_CALL(sblas_mem_advise(p, bytes, readOnly ? SBLAS_ADVISE_UNSET_READ_MOSTLY : SBLAS_ADVISE_UNSET_PREFERRED_LOCATION, gpu));
  if (prefetch_cpu) SBLAS_SAFE_CALL(sblas_mem_prefetch(p, bytes, -1));
}
inline void prefer_cpu(void* p, size_t bytes) {
  if (p == NULL || bytes == 0) return;
  SBLAS_SAFE_CALL(sblas_mem_advise(p, bytes, SBLAS_ADVISE_SET_PREFERRED_LOCATION, -1));
}
}  // namespace sblas_um_detail

/* ================================== CSR matrix (managed) ================================== */
template <typename IdxType, typename DataType>
class CsrSparseMatrix {
  static_assert(sizeof(IdxType) == 4, "the B200 build supports 32-bit indices (sblas_b200.h)");

 public:
  CsrSparseMatrix() { clear_(); }
  explicit CsrSparseMatrix(const char* filename) { /* matrix_um.h:419-456 */
    clear_();
    int32_t m = 0, n = 0, nnzA = 0, sym = 0;
    SBLAS_SAFE_CALL(sblas_mtx_info(filename, &m, &n, &nnzA, &sym));
    height = (IdxType)m;
    width = (IdxType)n;
    nnz = (IdxType)nnzA;
    SAFE_ALOC_MANAGED(csrRowPtr, get_row_ptr_size());
    SAFE_ALOC_MANAGED(csrColIdx, get_col_idx_size());
    SAFE_ALOC_MANAGED(csrVal, get_val_size());
    double* tmp = (double*)malloc(sizeof(double) * (size_t)std::max(nnzA, 1));
    SBLAS_SAFE_CALL(sblas_mtx_data(filename, (int32_t*)csrRowPtr, (int32_t*)csrColIdx, tmp));
    for (int32_t j = 0; j < nnzA; j++) csrVal[j] = (DataType)tmp[j];
    free(tmp);
    printf("input matrix A: ( %i, %i ) nnz = %i\n", m, n, nnzA);
  }
  /* Not in the reference: adopt host arrays (copied into managed memory), for synthetic matrices. */
  CsrSparseMatrix(IdxType h, IdxType w, const IdxType* rowptr, const IdxType* colidx, const DataType* val) {
    clear_();
    height = h;
    width = w;
    nnz = rowptr[h];
    SAFE_ALOC_MANAGED(csrRowPtr, get_row_ptr_size());
    SAFE_ALOC_MANAGED(csrColIdx, get_col_idx_size());
    SAFE_ALOC_MANAGED(csrVal, get_val_size());
    memcpy(csrRowPtr, rowptr, get_row_ptr_size());
    memcpy(csrColIdx, colidx, get_col_idx_size());
    memcpy(csrVal, val, get_val_size());
  }
  ~CsrSparseMatrix() {
    release_gpu_();
    SAFE_FREE_MANAGED(csrRowPtr);
    SAFE_FREE_MANAGED(csrColIdx);
    SAFE_FREE_MANAGED(csrVal);
  }
  CsrSparseMatrix(const CsrSparseMatrix&) = delete;
  CsrSparseMatrix& operator=(const CsrSparseMatrix&) = delete;

  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) { /* matrix_um.h:458-527 */
    sblas_range range__("sblas_um sync2gpu (partition + alias)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    release_gpu_();
    n_gpu = _n_gpu;
    policy = _policy;
    csrRowPtr_gpu = new IdxType*[n_gpu]();
    csrColIdx_gpu = new IdxType*[n_gpu]();
    csrVal_gpu = new DataType*[n_gpu]();
    plan_gpu = new sblasPlan_t[n_gpu]();
    if (policy == replicate) {
      for (unsigned i = 0; i < n_gpu; i++) {
        if (i == 0) { /* replica 0 is the base arrays themselves */
          csrRowPtr_gpu[i] = csrRowPtr;
          csrColIdx_gpu[i] = csrColIdx;
          csrVal_gpu[i] = csrVal;
        } else {
          SAFE_ALOC_MANAGED(csrRowPtr_gpu[i], get_row_ptr_size());
          SAFE_ALOC_MANAGED(csrColIdx_gpu[i], get_col_idx_size());
          SAFE_ALOC_MANAGED(csrVal_gpu[i], get_val_size());
          memcpy(csrRowPtr_gpu[i], csrRowPtr, get_row_ptr_size());
          memcpy(csrColIdx_gpu[i], csrColIdx, get_col_idx_size());
          memcpy(csrVal_gpu[i], csrVal, get_val_size());
        }
        SBLAS_SAFE_CALL(sblas_plan_create_on((int)i, (int32_t)height, (int64_t)nnz, (const int32_t*)csrRowPtr_gpu[i], &plan_gpu[i]));
        /* SpMV tile table + column-locality statistic now, and the column-panel mirror for gather-heavy inputs (a device
           copy owned by the plan; building it touches this replica's managed pages on GPU i) */
        SBLAS_SAFE_CALL(sblas_plan_prepare_spmv_on((int)i, sizeof(DataType) == 8, plan_gpu[i], (const int32_t*)csrColIdx_gpu[i],
                                                   csrVal_gpu[i], (int32_t)width));
      }
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
          SAFE_ALOC_MANAGED(csrRowPtr_gpu[i], get_gpu_row_ptr_size(i));
          SBLAS_SAFE_CALL(sblas_csr_segment_rowptr_i32((const int32_t*)csrRowPtr, avg_nnz, i,
                                                       (int32_t)starting_row_gpu[i], (int32_t)stoping_row_gpu[i],
                                                       (int32_t)nnz_gpu[i], (int32_t*)csrRowPtr_gpu[i]));
          csrColIdx_gpu[i] = csrColIdx + (size_t)i * (size_t)avg_nnz; /* aliases, matrix_um.h:517-518 */
          csrVal_gpu[i] = csrVal + (size_t)i * (size_t)avg_nnz;
          SBLAS_SAFE_CALL(sblas_plan_create_on((int)i, (int32_t)(get_gpu_row_ptr_num(i) - 1), (int64_t)nnz_gpu[i],
                                               (const int32_t*)csrRowPtr_gpu[i], &plan_gpu[i]));
          SBLAS_SAFE_CALL(sblas_plan_prepare_spmv_on((int)i, sizeof(DataType) == 8, plan_gpu[i], (const int32_t*)csrColIdx_gpu[i],
                                                     csrVal_gpu[i], (int32_t)width));
        }
        printf("gpu-%d,start-row:%d,stop-row:%d,num-rows:%ld,num-nnz:%d\n", i, (int)starting_row_gpu[i],
               (int)stoping_row_gpu[i], (long)get_gpu_row_ptr_num(i), (int)nnz_gpu[i]);
      }
    }
  }

  /* matrix_um.h:529-564 / :566-602 / :604-614 */
  void applyGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um applyGpuTuning");
    for (unsigned i = 0; i < n_gpu; i++) {
      sblas_um_detail::tune_range(csrRowPtr_gpu[i], gpu_rp_bytes_(i), readOnly, (int)i);
      sblas_um_detail::tune_range(csrColIdx_gpu[i], gpu_nz_(i) * sizeof(IdxType), readOnly, (int)i);
      sblas_um_detail::tune_range(csrVal_gpu[i], gpu_nz_(i) * sizeof(DataType), readOnly, (int)i);
    }
  }
  void removeGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um removeGpuTuning");
    for (unsigned i = 0; i < n_gpu; i++) {
      sblas_um_detail::untune_range(csrRowPtr_gpu[i], gpu_rp_bytes_(i), readOnly, (int)i, true);
      sblas_um_detail::untune_range(csrColIdx_gpu[i], gpu_nz_(i) * sizeof(IdxType), readOnly, (int)i, true);
      sblas_um_detail::untune_range(csrVal_gpu[i], gpu_nz_(i) * sizeof(DataType), readOnly, (int)i, true);
    }
  }
  void applyCpuTuning() {
    sblas_um_detail::prefer_cpu(csrRowPtr, get_row_ptr_size());
    sblas_um_detail::prefer_cpu(csrColIdx, get_col_idx_size());
    sblas_um_detail::prefer_cpu(csrVal, get_val_size());
  }

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

 public: /* matrix_um.h:651-671 */
  IdxType* csrRowPtr;
  IdxType* csrColIdx;
  DataType* csrVal;
  IdxType** csrRowPtr_gpu;
  IdxType** csrColIdx_gpu;
  DataType** csrVal_gpu;
  IdxType* nnz_gpu;
  IdxType* starting_row_gpu;
  IdxType* stoping_row_gpu;
  IdxType nnz;
  IdxType height;
  IdxType width;
  unsigned n_gpu;
  enum GpuSharePolicy policy;
  sblasPlan_t* plan_gpu; /* B200 build: per-GPU execution plans */

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
  size_t gpu_rp_bytes_(unsigned i) const { return policy == segment ? get_gpu_row_ptr_size(i) : get_row_ptr_size(); }
  size_t gpu_nz_(unsigned i) const { return policy == segment ? get_gpu_col_idx_num(i) : (size_t)nnz; }
  void release_gpu_() {
    if (plan_gpu != NULL) {
      for (unsigned i = 0; i < n_gpu; i++)
        if (plan_gpu[i] != NULL) sblas_plan_destroy(plan_gpu[i]);
      delete[] plan_gpu;
      plan_gpu = NULL;
    }
    if (csrRowPtr_gpu != NULL) {
      for (unsigned i = 0; i < n_gpu; i++) {
        /* owned: every segment row-pointer array, and replicas 1.. of a replicated matrix */
        const bool owned = (policy == segment) || i > 0;
        if (owned) SAFE_FREE_MANAGED(csrRowPtr_gpu[i]);
        if (policy == replicate && i > 0) {
          SAFE_FREE_MANAGED(csrColIdx_gpu[i]);
          SAFE_FREE_MANAGED(csrVal_gpu[i]);
        }
      }
      delete[] csrRowPtr_gpu;
      delete[] csrColIdx_gpu;
      delete[] csrVal_gpu;
      csrRowPtr_gpu = csrColIdx_gpu = NULL;
      csrVal_gpu = NULL;
    }
    SAFE_FREE_HOST(nnz_gpu);
    SAFE_FREE_HOST(starting_row_gpu);
    SAFE_FREE_HOST(stoping_row_gpu);
    n_gpu = 0;
    policy = none;
  }
};

/* ================================== dense matrix (managed) ================================== */
template <typename IdxType, typename DataType>
class DenseMatrix {
 public:
  DenseMatrix() : height(0), width(0), order(row_major) { init_(); }
  DenseMatrix(IdxType _height, IdxType _width, enum MajorOrder _order) : height(_height), width(_width), order(_order) {
    init_();
    SAFE_ALOC_MANAGED(val, get_mtx_size());
    srand(RAND_INIT_SEED);
    const size_t n = get_mtx_num();
    for (size_t i = 0; i < n; i++) val[i] = (DataType)rand0to1();
  }
  DenseMatrix(IdxType _height, IdxType _width, DataType _val, enum MajorOrder _order)
      : height(_height), width(_width), order(_order) {
    init_();
    SAFE_ALOC_MANAGED(val, get_mtx_size());
    srand(RAND_INIT_SEED);
    std::fill(val, val + get_mtx_num(), _val);
  }
  ~DenseMatrix() {
    release_gpu_();
    SAFE_FREE_MANAGED(val);
  }
  DenseMatrix(const DenseMatrix&) = delete;
  DenseMatrix& operator=(const DenseMatrix&) = delete;

  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) { /* matrix_um.h:760-791 */
    sblas_range range__("sblas_um sync2gpu (partition + alias)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    release_gpu_();
    n_gpu = _n_gpu;
    policy = _policy;
    const IdxType first_order = (order == row_major) ? height : width;
    const IdxType second_order = (order == row_major) ? width : height;
    val_gpu = new DataType*[n_gpu]();
    ld_gpu = new int64_t[n_gpu]();
    for (unsigned i = 0; i < n_gpu; i++) ld_gpu[i] = second_order;
    if (policy == replicate) {
      for (unsigned i = 0; i < n_gpu; i++) {
        if (i == 0) val_gpu[i] = val;
        else {
          SAFE_ALOC_MANAGED(val_gpu[i], get_mtx_size());
          memcpy(val_gpu[i], val, get_mtx_size());
        }
      }
    } else {
      SAFE_ALOC_HOST(dim_gpu, n_gpu * sizeof(IdxType));
      int32_t avg = 0;
      SBLAS_SAFE_CALL(sblas_dense_split_i32((int32_t)first_order, n_gpu, &avg, (int32_t*)dim_gpu));
      avg_val = (IdxType)avg;
      for (unsigned i = 0; i < n_gpu; i++)
        val_gpu[i] = dim_gpu[i] > 0 ? val + (size_t)i * (size_t)avg * (size_t)second_order : NULL; /* alias, :788 */
    }
  }
  void applyGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um applyGpuTuning"); /* matrix_um.h:793-814 */
    for (unsigned i = 0; i < n_gpu; i++) sblas_um_detail::tune_range(val_gpu[i], gpu_bytes_(i), readOnly, (int)i);
  }
  void removeGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um removeGpuTuning"); /* matrix_um.h:816-832 (no prefetch back) */
    for (unsigned i = 0; i < n_gpu; i++) sblas_um_detail::untune_range(val_gpu[i], gpu_bytes_(i), readOnly, (int)i, false);
  }
  void applyCpuTuning() { sblas_um_detail::prefer_cpu(val, get_mtx_size()); }

  /* matrix_um.h:870-890 — nothing to copy; optionally pull GPU i's range to the CPU, then wait */
  void sync2cpu(unsigned i_gpu, bool tuning = false) {
    assert(val_gpu != NULL);
    assert(i_gpu < n_gpu);
    if (tuning) SBLAS_SAFE_CALL(sblas_mem_prefetch(val_gpu[i_gpu], gpu_bytes_(i_gpu), -1));
    sblas_sync_gpus(n_gpu);
  }
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
  void plusDenseMatrixGPU(DenseMatrix const& dm, DataType alpha, DataType beta) { /* matrix_um.h:891-904 */
    if (n_gpu == 0 || policy == none) return;
    for (unsigned i = 0; i < n_gpu; i++)
      if (val_gpu[i] != NULL)
        SBLAS_SAFE_CALL(sblas_axpby_on((int)i, sizeof(DataType) == 8, val_gpu[i], dm.val_gpu[i], (double)alpha,
                                       (double)beta, (int64_t)(gpu_bytes_(i) / sizeof(DataType))));
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

 public: /* matrix_um.h:915-926 */
  IdxType height;
  IdxType width;
  DataType* val;
  DataType** val_gpu;
  IdxType* dim_gpu;
  unsigned n_gpu;
  enum GpuSharePolicy policy;
  enum MajorOrder order;
  int32_t dev_layout; /* always the host layout here: the device "copies" alias the host array */
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
    dev_layout = SBLAS_COL_MAJOR;
    avg_val = 0;
  }
  size_t gpu_bytes_(unsigned i) const {
    if (policy != segment) return get_mtx_size();
    const IdxType second_order = (order == row_major) ? width : height;
    return dim_gpu[i] > 0 ? (size_t)second_order * (size_t)dim_gpu[i] * sizeof(DataType) : 0;
  }
  void release_gpu_() {
    if (val_gpu != NULL) {
      if (policy == replicate)
        for (unsigned i = 1; i < n_gpu; i++) SAFE_FREE_MANAGED(val_gpu[i]); /* replica 0 and slices alias val */
      delete[] val_gpu;
      val_gpu = NULL;
    }
    delete[] ld_gpu;
    ld_gpu = NULL;
    SAFE_FREE_HOST(dim_gpu);
    n_gpu = 0;
    policy = none;
  }
};

/* ================================== dense vector (managed) ================================== */
template <typename IdxType, typename DataType>
class DenseVector {
 public:
  DenseVector() : length(0), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {}
  explicit DenseVector(IdxType _length) : length(_length), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {
    SAFE_ALOC_MANAGED(val, get_vec_size());
    srand(RAND_INIT_SEED);
    for (size_t i = 0; i < get_vec_length(); i++) val[i] = (DataType)rand0to1();
  }
  DenseVector(IdxType _length, DataType _val) : length(_length), val(NULL), val_gpu(NULL), n_gpu(0), policy(none) {
    SAFE_ALOC_MANAGED(val, get_vec_size());
    srand(RAND_INIT_SEED);
    std::fill(val, val + get_vec_length(), _val);
  }
  /* matrix_um.h:958-972 — the copy gets its own replicas; its base array is its replica 0 */
  DenseVector(const DenseVector& dv) : length(dv.length), val(NULL), val_gpu(NULL), n_gpu(dv.n_gpu), policy(dv.policy) {
    SAFE_ALOC_MANAGED(val, get_vec_size());
    if (n_gpu != 0 && policy != none) {
      val_gpu = new DataType*[n_gpu]();
      for (unsigned i = 0; i < n_gpu; i++) {
        if (i == 0) val_gpu[i] = val; else SAFE_ALOC_MANAGED(val_gpu[i], get_vec_size());
        SBLAS_SAFE_CALL(sblas_copy_d2d((int)i, val_gpu[i], dv.val_gpu[i], get_vec_size()));
      }
    } else {
      memcpy(val, dv.val, get_vec_size());
    }
  }
  DenseVector& operator=(const DenseVector&) = delete;
  ~DenseVector() {
    release_gpu_();
    SAFE_FREE_MANAGED(val);
  }
  void sync2gpu(unsigned _n_gpu, enum GpuSharePolicy _policy) { /* matrix_um.h:973-999 */
    sblas_range range__("sblas_um sync2gpu (partition + alias)");
    assert(_n_gpu != 0);
    assert(_policy != none);
    assert(_policy != segment);
    release_gpu_();
    n_gpu = _n_gpu;
    policy = _policy;
    val_gpu = new DataType*[n_gpu]();
    for (unsigned i = 0; i < n_gpu; i++) {
      if (i == 0) val_gpu[i] = val;
      else {
        SAFE_ALOC_MANAGED(val_gpu[i], get_vec_size());
        memcpy(val_gpu[i], val, get_vec_size());
      }
    }
  }
  void applyGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um applyGpuTuning");
    for (unsigned i = 0; i < n_gpu; i++) sblas_um_detail::tune_range(val_gpu[i], get_vec_size(), readOnly, (int)i);
  }
  void removeGpuTuning(bool readOnly = true) { sblas_range range__("sblas_um removeGpuTuning");
    for (unsigned i = 0; i < n_gpu; i++) sblas_um_detail::untune_range(val_gpu[i], get_vec_size(), readOnly, (int)i, false);
  }
  void applyCpuTuning() { sblas_um_detail::prefer_cpu(val, get_vec_size()); }
  void sync2cpu(unsigned i_gpu, bool tuning = false) { /* matrix_um.h:1031-1045 */
    assert(i_gpu < n_gpu);
    assert(val_gpu != NULL);
    if (tuning) SBLAS_SAFE_CALL(sblas_mem_prefetch(val_gpu[i_gpu], get_vec_size(), -1));
    sblas_sync_gpus(n_gpu);
  }
  void plusDenseVectorGPU(DenseVector const& dv, DataType alpha, DataType beta) {
    if (n_gpu == 0 || policy == none) return;
    for (unsigned i = 0; i < n_gpu; i++)
      SBLAS_SAFE_CALL(sblas_axpby_on((int)i, sizeof(DataType) == 8, val_gpu[i], dv.val_gpu[i], (double)alpha,
                                     (double)beta, (int64_t)get_vec_length()));
    sblas_sync_gpus(n_gpu);
  }
  size_t get_vec_size() const { return (size_t)length * sizeof(DataType); }
  size_t get_vec_length() const { return (size_t)length; }

 public:
  IdxType length;
  DataType* val;
  DataType** val_gpu;
  unsigned n_gpu;
  enum GpuSharePolicy policy;

 private:
  void release_gpu_() {
    if (val_gpu != NULL) {
      for (unsigned i = 1; i < n_gpu; i++) SAFE_FREE_MANAGED(val_gpu[i]);
      delete[] val_gpu;
      val_gpu = NULL;
    }
    n_gpu = 0;
    policy = none;
  }
};

#endif /* MATRIX_UM_H */
