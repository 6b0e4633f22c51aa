/* include/sblas_um.h — umbrella header of the managed-memory variant (sblas_um.h:18-19). */
#ifndef SBLAS_UM_H
#define SBLAS_UM_H
#ifndef SBLAS_UM
#define SBLAS_UM 1
#endif
#include "sblas.h"
#endif /* SBLAS_UM_H */
