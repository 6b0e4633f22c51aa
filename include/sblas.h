/* include/sblas.h — umbrella header of the B200 build (the reference's sblas.h:18-19 pulls in
 * spmm.h and spmv.h the same way).  Link with -lsblas_b200 (s-blas_b200/lib). */
#ifndef SBLAS_H
#define SBLAS_H

#include <iostream>

#include "kernel.h"
#include "spmm.h" /* pulls in matrix.h, or matrix_um.h under sblas_um.h */
#include "spmv.h"

using std::cerr; /* the reference's headers put these names in scope for its drivers */
using std::cout;
using std::endl;

#endif /* SBLAS_H */
