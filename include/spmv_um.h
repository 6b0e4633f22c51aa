/* include/spmv_um.h — managed-memory SpMV entry points (drop-in for the reference's spmv_um.h):
 * the same templates as spmv.h over the classes of matrix_um.h, with the trailing
 * `bool tuning = false` argument of spmv_um.h:39. */
#ifndef SPMV_UM_H
#define SPMV_UM_H
#ifndef SBLAS_UM
#define SBLAS_UM 1
#endif
#include "spmv.h"
#endif /* SPMV_UM_H */
