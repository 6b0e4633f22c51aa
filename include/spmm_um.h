/* include/spmm_um.h — managed-memory SpMM entry points (drop-in for the reference's spmm_um.h):
 * the same templates as spmm.h over the classes of matrix_um.h, with the trailing
 * `bool tuning = false` argument of spmm_um.h:87 / :167. */
#ifndef SPMM_UM_H
#define SPMM_UM_H
#ifndef SBLAS_UM
#define SBLAS_UM 1
#endif
#include "spmm.h"
#endif /* SPMM_UM_H */
