/* include/kernel.h — the element-wise utility kernels of the reference's kernel.h, B200 build.
 *
 * The reference defines two __global__ templates and launches them with <<<grid, 256>>> from
 * matrix.h:620 / :721.  Here the device code lives in libsblas_b200.so (s-blas_b200/csrc/
 * dense.cu: 128-bit vectorised, grid sized to the SM count) and these are host functions that
 * enqueue it on logical GPU `gpu`'s default stream; the SpMV/SpMM ops no longer call them at
 * all because the alpha/beta update is fused into their epilogues. */
#ifndef KERNEL_H
#define KERNEL_H

#include "utility.h"

/* kernel.h:28-37 — vec0[i] = vec0[i]*beta + vec1[i]*alpha */
template <typename IdxType, typename DataType>
inline void denseVector_plusEqual_denseVector(int gpu, DataType* vec0, const DataType* vec1, DataType alpha,
                                              DataType beta, IdxType length) {
  SBLAS_SAFE_CALL(sblas_axpby_on(gpu, sizeof(DataType) == 8, vec0, vec1, (double)alpha, (double)beta, (int64_t)length));
}

/* kernel.h:19-25 — vec[i] = vec[i]*beta + val (never launched by the reference, and indexed
 * with blockIdx.x*gridDim.x there, kernel.h:22; the B200 kernel uses the intended index) */
template <typename IdxType, typename DataType>
inline void denseVector_plusEqual_scalar(int gpu, DataType* vec, DataType val, DataType beta, IdxType length) {
  SBLAS_SAFE_CALL(sblas_scal_add_on(gpu, sizeof(DataType) == 8, vec, (double)val, (double)beta, (int64_t)length));
}

#endif /* KERNEL_H */
