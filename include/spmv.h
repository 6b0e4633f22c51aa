/* include/spmv.h — CSR SpMV entry points, B200 build (drop-in for the reference's spmv.h).
 *
 *   sblas_spmv_csr_cpu   spmv.h:16-32   single-threaded CPU statement, kept as the checker the
 *                                       reference's drivers print "Validation" from; no op
 *                                       of this build calls it
 *   sblas_spmv_csr_v1    spmv.h:36-139  y = beta*y + alpha*A*x on n_gpu GPUs
 *
 * The reference's v1 creates NCCL communicators and a cuSPARSE handle per call, uploads a
 * zero vector to every GPU, runs cusparseSpMV on each nnz segment, all-reduces the whole
 * vector and finishes with an axpby pass.  Here v1 is ONE call into libsblas_b200.so
 * (sblas_mgpu_spmv_csr_v1_*): each GPU runs a hand-written sm_100a kernel over its segment
 * with the alpha/beta epilogue fused, stores finished rows straight into every GPU's replica
 * of y over NVLink, and the <= n_gpu-1 rows shared by neighbouring segments are completed from
 * per-row partial sums.  Pre- and post-conditions are the reference's: A segment-distributed,
 * x and y replicated; on return every replica of y is updated (matrix.h:714-726). */
#ifndef SPMV_H
#define SPMV_H

#ifdef SBLAS_UM /* set by spmv_um.h / spmm_um.h: managed-memory classes, extra `tuning` argument */
#include "matrix_um.h"
#define SBLAS_TUNING_PARAM , bool tuning = false
#define SBLAS_TUNING_USE (void)tuning
#else
#include "matrix.h"
#define SBLAS_TUNING_PARAM
#define SBLAS_TUNING_USE
#endif

template <typename IdxType, typename DataType>
void sblas_spmv_csr_cpu(CsrSparseMatrix<IdxType, DataType>* pA, DenseVector<IdxType, DataType>* pB,
                        DenseVector<IdxType, DataType>* pC, DataType alpha, DataType beta) {
  assert(pA->width == pB->length);
  assert(pB->length == pC->length); /* spmv.h:21: the checker only takes square A */
  const IdxType* rp = pA->csrRowPtr;
  for (IdxType r = 0; r < pA->height; r++) {
    DataType acc = 0; /* left to right in storage order, separately rounded mul and add */
    for (IdxType j = rp[r]; j < rp[r + 1]; j++) acc += pA->csrVal[j] * pB->val[pA->csrColIdx[j]];
    pC->val[r] = beta * pC->val[r] + alpha * acc;
  }
}

template <typename IdxType, typename DataType>
void sblas_spmv_csr_v1(CsrSparseMatrix<IdxType, DataType>* pA, DenseVector<IdxType, DataType>* pB,
                       DenseVector<IdxType, DataType>* pC, DataType alpha, DataType beta, unsigned n_gpu SBLAS_TUNING_PARAM) {
  sblas_range range__("sblas_spmv_csr_v1");
  SBLAS_TUNING_USE; /* the reference only uses it to advise its C_copy temporary (spmv_um.h:51-53), which no longer exists */
  assert(pA->width == pB->length);  /* spmv.h:40 */
  assert(pA->height == pC->length); /* spmv.h:41 */
  if (pA->policy != segment || pA->n_gpu < n_gpu || pB->n_gpu < n_gpu || pC->n_gpu < n_gpu) {
    fprintf(stderr, "SBLAS_SPMV_CSR_V1: A must be sync2gpu(n_gpu, segment), x and y sync2gpu(n_gpu, replicate)\n");
    exit(-1);
  }
  SBLAS_SAFE_CALL(sblas_abi<DataType>::spmv_v1(
      n_gpu, (int32_t)pA->height, (int32_t)pA->width, (const int32_t* const*)pA->csrRowPtr_gpu,
      (const int32_t* const*)pA->csrColIdx_gpu, pA->csrVal_gpu, (const int32_t*)pA->nnz_gpu,
      (const int32_t*)pA->starting_row_gpu, (const int32_t*)pA->stoping_row_gpu, pB->val_gpu, pC->val_gpu, alpha,
      beta, pA->plan_gpu));
}

#endif /* SPMV_H */
