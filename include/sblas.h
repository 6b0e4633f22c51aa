/* include/sblas.h — umbrella header of the B200 build: one include gives a driver the two ops
 * (spmv.h, spmm.h), the data model they pull in (matrix.h, or matrix_um.h under sblas_um.h), the
 * utility kernels and the stream names the reference's drivers use unqualified (its own headers
 * put cout / cerr / endl in scope).  The reference's sblas.h:18-19 plays the same role.
 * Link with -lsblas_b200 (s-blas_b200/lib). */
#pragma once
#if !defined(SBLAS_H)
#define SBLAS_H 1 /* the reference's include guard, for code that tests it */
#endif

#include <iostream>
using std::cerr;
using std::cout;
using std::endl;

#include "kernel.h"
#include "spmv.h"
#include "spmm.h"
