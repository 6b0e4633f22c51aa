/* include/spmm.h — CSR SpMM entry points, B200 build (drop-in for the reference's spmm.h).
 *
 *   sblas_spmm_csr_cpu   spmm.h:30-70    CPU statement (checker for the drivers)
 *   sblas_spmm_csr_v1    spmm.h:84-161   method 1: A replicated, B and C split by columns
 *   sblas_spmm_csr_v2    spmm.h:164-284  method 2: A nnz-segmented, B and C replicated
 *
 * C = beta*C + alpha*A*B with B, C col_major (row_major operands are rejected exactly like
 * spmm.h:91-98).  Each op is one call into libsblas_b200.so; there is no cuSPARSE, no NCCL
 * all-reduce over the whole of C, no zero-filled C_copy and no separate axpby pass.  v1 keeps
 * the reference's side effect of refreshing the host copy of every C slice (spmm.h:151); v2
 * leaves the result in every device replica, the caller does C.sync2cpu(0) (spmm_test.cu:84). */
#ifndef SPMM_H
#define SPMM_H

#include <iostream>

#ifdef SBLAS_UM
#include "matrix_um.h"
#else
#include "matrix.h"
#endif
#ifndef SBLAS_TUNING_USE
#ifdef SBLAS_UM
#define SBLAS_TUNING_PARAM , bool tuning = false
#define SBLAS_TUNING_USE (void)tuning
#else
#define SBLAS_TUNING_PARAM
#define SBLAS_TUNING_USE
#endif
#endif

template <typename IdxType, typename DataType>
void sblas_spmm_csr_cpu(CsrSparseMatrix<IdxType, DataType>* pA, DenseMatrix<IdxType, DataType>* pB,
                        DenseMatrix<IdxType, DataType>* pC, DataType alpha, DataType beta) {
  assert(pA->width == pB->height);
  assert(pA->height == pC->height);
  assert(pB->width == pC->width);
  if (pB->order != col_major || pC->order != col_major) {
    /* spmm.h:40-54 has a row-major-C branch that mixes layouts; only the
     * col-major statement (spmm.h:55-68) is defined behaviour, so that is all this build states */
    std::cerr << "sblas_spmm_csr_cpu: B and C should be in column major!" << std::endl;
    exit(-1);
  }
  const size_t hB = (size_t)pB->height, hC = (size_t)pC->height;
  for (IdxType r = 0; r < pA->height; r++)
    for (IdxType n = 0; n < pB->width; n++) {
      const DataType* bcol = pB->val + (size_t)n * hB;
      DataType acc = 0;
      for (IdxType j = pA->csrRowPtr[r]; j < pA->csrRowPtr[r + 1]; j++) acc += pA->csrVal[j] * bcol[pA->csrColIdx[j]];
      DataType* c = pC->val + (size_t)n * hC + r;
      *c = beta * (*c) + alpha * acc;
    }
}

template <typename IdxType, typename DataType>
inline void sblas_require_col_major_(const DenseMatrix<IdxType, DataType>* pB, const DenseMatrix<IdxType, DataType>* pC,
                                     const char* who) {
  if (pB->order == row_major) { /* spmm.h:91-94 */
    std::cerr << who << ": B should be in column major!" << std::endl;
    exit(-1);
  }
  if (pC->order == row_major) { /* spmm.h:95-98 */
    std::cerr << who << ": C should be in column major!" << std::endl;
    exit(-1);
  }
  if (pB->dev_layout != pC->dev_layout) {
    std::cerr << who << ": B and C must use the same device layout" << std::endl;
    exit(-1);
  }
}

/* method 1 */
template <typename IdxType, typename DataType>
void sblas_spmm_csr_v1(CsrSparseMatrix<IdxType, DataType>* pA, DenseMatrix<IdxType, DataType>* pB,
                       DenseMatrix<IdxType, DataType>* pC, DataType alpha, DataType beta, unsigned n_gpu SBLAS_TUNING_PARAM) {
  sblas_range range__("sblas_spmm_csr_v1");
  SBLAS_TUNING_USE; /* the reference only uses it to advise its C_copy temporary (spmv_um.h:51-53), which no longer exists */
  assert(pA->width == pB->height);
  assert(pA->height == pC->height);
  assert(pB->width == pC->width);
  sblas_require_col_major_(pB, pC, "SBLAS_SPMM_CSR_V1");
  std::cout << "sblas_spmm_csr_v1 ready to start" << std::endl; /* spmm.h:99 */
  if (pA->policy != replicate || pB->policy != segment || pC->policy != segment || pA->n_gpu < n_gpu ||
      pB->n_gpu < n_gpu || pC->n_gpu < n_gpu) {
    std::cerr << "SBLAS_SPMM_CSR_V1: A must be replicated, B and C segmented over n_gpu GPUs" << std::endl;
    exit(-1);
  }
  SBLAS_SAFE_CALL(sblas_abi<DataType>::spmm_v1(
      n_gpu, (int32_t)pA->height, (int32_t)pA->width, (int32_t)pA->nnz, (const int32_t* const*)pA->csrRowPtr_gpu,
      (const int32_t* const*)pA->csrColIdx_gpu, pA->csrVal_gpu, (const int32_t*)pB->dim_gpu, pB->dev_layout,
      pB->val_gpu, pB->ld_gpu, pC->val_gpu, pC->ld_gpu, alpha, beta, pA->plan_gpu));
#ifdef SBLAS_UM
  for (unsigned i = 0; i < n_gpu; i++) pC->sync2cpu(i); /* spmm.h:151; managed memory: nothing is copied */
#else
  pC->sync2cpu_all(n_gpu); /* spmm.h:151 for every GPU, the slices' downloads in flight together */
#endif
}

/* method 2 */
template <typename IdxType, typename DataType>
void sblas_spmm_csr_v2(CsrSparseMatrix<IdxType, DataType>* pA, DenseMatrix<IdxType, DataType>* pB,
                       DenseMatrix<IdxType, DataType>* pC, DataType alpha, DataType beta, unsigned n_gpu SBLAS_TUNING_PARAM) {
  sblas_range range__("sblas_spmm_csr_v2");
  SBLAS_TUNING_USE; /* the reference only uses it to advise its C_copy temporary (spmv_um.h:51-53), which no longer exists */
  assert(pA->width == pB->height);
  assert(pA->height == pC->height);
  assert(pB->width == pC->width);
  sblas_require_col_major_(pB, pC, "SBLAS_SPMM_CSR_V2");
  if (pA->policy != segment || pB->policy != replicate || pC->policy != replicate || pA->n_gpu < n_gpu ||
      pB->n_gpu < n_gpu || pC->n_gpu < n_gpu) {
    std::cerr << "SBLAS_SPMM_CSR_V2: A must be segmented, B and C replicated over n_gpu GPUs" << std::endl;
    exit(-1);
  }
  SBLAS_SAFE_CALL(sblas_abi<DataType>::spmm_v2(
      n_gpu, (int32_t)pA->height, (int32_t)pA->width, (const int32_t* const*)pA->csrRowPtr_gpu,
      (const int32_t* const*)pA->csrColIdx_gpu, pA->csrVal_gpu, (const int32_t*)pA->nnz_gpu,
      (const int32_t*)pA->starting_row_gpu, (const int32_t*)pA->stoping_row_gpu, (int32_t)pB->width, pB->dev_layout,
      pB->val_gpu, pB->ld_gpu[0], pC->val_gpu, pC->ld_gpu[0], alpha, beta, pA->plan_gpu));
}

#endif /* SPMM_H */
