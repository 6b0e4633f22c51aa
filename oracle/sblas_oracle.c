/* oracle/sblas_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's (tartarughina/S-BLAS) CSR SpMV / SpMM
 * verification path and of the integer partition arithmetic around it.  Each
 * function cites the reference file:line it follows.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this; the product (s-blas_b200/) never does.
 *
 * PARITY PINNED: tests/test_oracle.py checks every function here against
 *   (a) the golden vectors in tests/golden/ (generated from the reference's own
 *       code by oracle/gen_golden.py through oracle/_ref/libsblas_ref.so), and
 *   (b) when oracle/_ref/libsblas_ref.so is present, the reference itself on seeded
 *       random inputs (bit-for-bit).
 *
 * Build: see oracle/Makefile  (gcc -O3 -fopenmp -ffp-contract=off -shared -fPIC).
 * -ffp-contract=off keeps "sum += a*b" as a rounded multiply followed by a rounded
 * add, which is what the reference's x86-64 build executes (no FMA in its flags,
 * Makefile:23), so the restatement is bit-identical to it.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* config.h:21,23 */
#define ERROR_BAR (1e-3)
#define RAND_INIT_SEED 211

/* ---------------------------------------------------------------- SpMV ---- */
/* spmv.h:16-32: y[i] = beta*y[i] + alpha * sum_j val[j]*x[col[j]], j ascending,
 * sum starts at 0, single thread. */
#define DEF_SPMV(NAME, T)                                                              \
  ORACLE_API void NAME(int h, const int* rp, const int* ci, const T* v, const T* x,   \
                       T* y, T alpha, T beta) {                                        \
    for (int i = 0; i < h; i++) {                                                      \
      T sum = 0;                                                                       \
      for (int j = rp[i]; j < rp[i + 1]; j++) sum += v[j] * x[ci[j]];                  \
      y[i] = beta * y[i] + alpha * sum;                                                \
    }                                                                                  \
  }
DEF_SPMV(oracle_spmv_csr_f64, double)
DEF_SPMV(oracle_spmv_csr_f32, float)

/* Same loop, rows [r0,r1) only, OpenMP row-parallel (identical per-row order, so
 * bit-identical to the serial one).  Used as the timed all-cores CPU baseline. */
#define DEF_SPMV_OMP(NAME, T)                                                          \
  ORACLE_API void NAME(int r0, int r1, const int* rp, const int* ci, const T* v,      \
                       const T* x, T* y, T alpha, T beta, int nthreads) {              \
    _Pragma("omp parallel for schedule(static) num_threads(nthreads)")                 \
    for (int i = r0; i < r1; i++) {                                                    \
      T sum = 0;                                                                       \
      for (int j = rp[i]; j < rp[i + 1]; j++) sum += v[j] * x[ci[j]];                  \
      y[i] = beta * y[i] + alpha * sum;                                                \
    }                                                                                  \
  }
DEF_SPMV_OMP(oracle_spmv_csr_rows_omp_f64, double)
DEF_SPMV_OMP(oracle_spmv_csr_rows_omp_f32, float)

/* ---------------------------------------------------------------- SpMM ---- */
/* spmm.h:55-68 (col-major B [w x k] with leading dim Bh, col-major C [h x k] with
 * leading dim Ch): loop order i, n, j; sum starts at 0. */
#define DEF_SPMM(NAME, T)                                                              \
  ORACLE_API void NAME(int h, const int* rp, const int* ci, const T* v, int k,        \
                       const T* B, long Bh, T* C, long Ch, T alpha, T beta) {          \
    for (int i = 0; i < h; i++) {                                                      \
      for (int n = 0; n < k; n++) {                                                    \
        T sum = 0;                                                                     \
        for (int j = rp[i]; j < rp[i + 1]; j++) sum += v[j] * B[n * Bh + ci[j]];       \
        C[n * Ch + i] = beta * C[n * Ch + i] + alpha * sum;                            \
      }                                                                                \
    }                                                                                  \
  }
DEF_SPMM(oracle_spmm_csr_f64, double)
DEF_SPMM(oracle_spmm_csr_f32, float)

#define DEF_SPMM_OMP(NAME, T)                                                          \
  ORACLE_API void NAME(int r0, int r1, const int* rp, const int* ci, const T* v,      \
                       int k, const T* B, long Bh, T* C, long Ch, T alpha, T beta,     \
                       int nthreads) {                                                 \
    _Pragma("omp parallel for schedule(static) num_threads(nthreads)")                 \
    for (int i = r0; i < r1; i++) {                                                    \
      for (int n = 0; n < k; n++) {                                                    \
        T sum = 0;                                                                     \
        for (int j = rp[i]; j < rp[i + 1]; j++) sum += v[j] * B[n * Bh + ci[j]];       \
        C[n * Ch + i] = beta * C[n * Ch + i] + alpha * sum;                            \
      }                                                                                \
    }                                                                                  \
  }
DEF_SPMM_OMP(oracle_spmm_csr_rows_omp_f64, double)
DEF_SPMM_OMP(oracle_spmm_csr_rows_omp_f32, float)

/* |A||x| row bounds used by the north-star tolerance (per-row error relative to
 * (|A||x|)_i); not in the reference — it is the yardstick, not the oracle. */
ORACLE_API void oracle_absrow_spmv_f64(int h, const int* rp, const int* ci, const double* v,
                                       const double* x, double* out) {
  for (int i = 0; i < h; i++) {
    double s = 0;
    for (int j = rp[i]; j < rp[i + 1]; j++) s += fabs(v[j]) * fabs(x[ci[j]]);
    out[i] = s;
  }
}

/* kernel.h:28-37: v0 = v0*beta + v1*alpha */
ORACLE_API void oracle_axpby_f64(double* v0, const double* v1, double alpha, double beta, long n) {
  for (long i = 0; i < n; i++) v0[i] = v0[i] * beta + v1[i] * alpha;
}
ORACLE_API void oracle_axpby_f32(float* v0, const float* v1, float alpha, float beta, long n) {
  for (long i = 0; i < n; i++) v0[i] = v0[i] * beta + v1[i] * alpha;
}

/* utility.h:182-191: absolute difference > ERROR_BAR fails. */
ORACLE_API int oracle_check_equal_f64(const double* x, const double* y, size_t m) {
  int ok = 1;
  for (size_t i = 0; i < m; i++)
    if (fabs(x[i] - y[i]) > ERROR_BAR) ok = 0;
  return ok;
}
ORACLE_API int oracle_check_equal_f32(const float* x, const float* y, size_t m) {
  int ok = 1;
  for (size_t i = 0; i < m; i++)
    if (fabsf(x[i] - y[i]) > ERROR_BAR) ok = 0;
  return ok;
}

/* ----------------------------------------------------------- partition ---- */
/* utility.h:293-300: first row i with rp[i] <= idx < rp[i+1] (linear scan, skips
 * empty rows), -1 when none. */
ORACLE_API int oracle_find_row(const int* rp, int h, int idx) {
  for (int i = 0; i < h; i++)
    if (rp[i] <= idx && idx < rp[i + 1]) return i;
  return -1;
}

/* matrix.h:360-375, CsrSparseMatrix::sync2gpu(segment).
 * float_ceil != 0 reproduces the reference's `ceil((float)nnz / n_gpu)` (float32
 * cast, matrix.h:360) literally, including its loss of trailing nnz when
 * nnz > 2^24.  float_ceil == 0 is the integer ceil-div the product
 * uses; the two agree whenever the float division is exact.
 * rowptr_out (may be NULL): the n_gpu local row-pointer arrays back to back,
 * rowptr_off[i] = start of GPU i's array (length stop-start+2, matrix.h:398-404). */
ORACLE_API int oracle_partition_nnz(const int* rp, int height, int nnz, unsigned n_gpu,
                                    int float_ceil, int* avg_out, int* nnz_gpu, int* start_row,
                                    int* stop_row, int* rowptr_out, long* rowptr_off) {
  int avg_nnz;
  if (float_ceil)
    avg_nnz = (int)ceil((float)nnz / n_gpu);
  else
    avg_nnz = (int)(((long)nnz + n_gpu - 1) / n_gpu);
  *avg_out = avg_nnz;
  long off = 0;
  for (unsigned i = 0; i < n_gpu; i++) {
    int s_nnz = (int)(i * avg_nnz);                                   /* :363 */
    int hi = (int)((i + 1) * avg_nnz);
    int e_nnz = (hi < nnz ? hi : nnz) - 1;                            /* :364 */
    nnz_gpu[i] = e_nnz - s_nnz + 1;                                   /* :365 */
    start_row[i] = oracle_find_row(rp, height, s_nnz);                /* :366 */
    stop_row[i] = oracle_find_row(rp, height, e_nnz);                 /* :368 */
    long num = (long)stop_row[i] - start_row[i] + 2;                  /* :401 */
    rowptr_off[i] = off;
    if (rowptr_out) {
      int* f = rowptr_out + off;
      f[0] = 0;                                                       /* :372 */
      for (long k = 1; k < num - 1; k++)
        f[k] = rp[start_row[i] + k] - (int)(i * avg_nnz);             /* :374 */
      f[num - 1] = nnz_gpu[i];                                        /* :375 */
    }
    off += num;
  }
  rowptr_off[n_gpu] = off;
  return 0;
}

/* matrix.h:557-567, DenseMatrix::sync2gpu(segment): split of the leading
 * ("first order") dimension; dim may be <= 0 on trailing GPUs. */
ORACLE_API void oracle_dense_split(int first_order, unsigned n_gpu, int* avg_out, int* dim_gpu) {
  int avg_val = (int)ceil((double)first_order / n_gpu);
  *avg_out = avg_val;
  for (unsigned i = 0; i < n_gpu; i++) {
    int hi = (int)((i + 1) * avg_val);
    dim_gpu[i] = (hi < first_order ? hi : first_order) - (int)(i * avg_val);
  }
}

/* matrix.h:522-524 + utility.h:197: srand(211); val[i] = rand()/RAND_MAX (glibc). */
ORACLE_API void oracle_rand_fill_f64(double* p, long n) {
  srand(RAND_INIT_SEED);
  for (long i = 0; i < n; i++) p[i] = (double)rand() / (double)RAND_MAX;
}
ORACLE_API void oracle_rand_fill_f32(float* p, long n) {
  srand(RAND_INIT_SEED);
  for (long i = 0; i < n; i++) p[i] = (float)((double)rand() / (double)RAND_MAX);
}

/* ------------------------------------------------------ Matrix Market ---- */
/* mmio.h:254-337 (banner), :339-367 (size line), mmio_highlevel.h:7-127 (info) and
 * :130-281 (data).  Coordinate format only (the reference's fscanf patterns only
 * parse coordinate entries).  Semantics restated:
 *   - pattern -> value 1.0; integer -> value read with %d; real -> %lg; complex ->
 *     real part kept (mmio_highlevel.h:196-206);
 *   - symmetric OR hermitian (skew-symmetric is NOT expanded: mm_is_symmetric tests
 *     'S' only, mmio.h) -> off-diagonal entries mirrored;
 *   - rows filled by counting sort in FILE ORDER, mirrored entry placed right after
 *     its original in its own row; columns are NOT sorted within a row.           */
typedef struct { int pattern, real, cplx, integer, sym; } mm_kind;

static int mm_banner(FILE* f, mm_kind* k) {
  char line[1100], banner[64], mtx[64], crd[64], dtype[64], symm[64];
  if (!fgets(line, sizeof line, f)) return -1;
  if (sscanf(line, "%63s %63s %63s %63s %63s", banner, mtx, crd, dtype, symm) != 5) return -1;
  for (char* p = mtx; *p; p++) *p = (char)tolower(*p);
  for (char* p = crd; *p; p++) *p = (char)tolower(*p);
  for (char* p = dtype; *p; p++) *p = (char)tolower(*p);
  for (char* p = symm; *p; p++) *p = (char)tolower(*p);
  if (strncmp(banner, "%%MatrixMarket", 14) != 0) return -1;
  if (strcmp(mtx, "matrix") != 0) return -1;
  if (strcmp(crd, "coordinate") != 0 && strcmp(crd, "array") != 0) return -1;
  memset(k, 0, sizeof *k);
  if (!strcmp(dtype, "real")) k->real = 1;
  else if (!strcmp(dtype, "complex")) k->cplx = 1;
  else if (!strcmp(dtype, "pattern")) k->pattern = 1;
  else if (!strcmp(dtype, "integer")) k->integer = 1;
  else return -1;
  if (!strcmp(symm, "symmetric") || !strcmp(symm, "hermitian")) k->sym = 1;
  else if (strcmp(symm, "general") && strcmp(symm, "skew-symmetric")) return -1;
  return 0;
}
static int mm_size(FILE* f, int* m, int* n, int* nz) {
  char line[1100];
  do {
    if (!fgets(line, sizeof line, f)) return -1;
  } while (line[0] == '%');
  if (sscanf(line, "%d %d %d", m, n, nz) == 3) return 0;
  int got;
  do {
    got = fscanf(f, "%d %d %d", m, n, nz);
    if (got == EOF) return -1;
  } while (got != 3);
  return 0;
}
static int mm_load(const char* fn, mm_kind* k, int* m, int* n, int* nz, int** ri, int** cj,
                   double** vv) {
  FILE* f = fopen(fn, "r");
  if (!f) return -1;
  if (mm_banner(f, k)) { fclose(f); return -2; }
  if (mm_size(f, m, n, nz)) { fclose(f); return -4; }
  *ri = (int*)malloc(sizeof(int) * (size_t)(*nz > 0 ? *nz : 1));
  *cj = (int*)malloc(sizeof(int) * (size_t)(*nz > 0 ? *nz : 1));
  *vv = (double*)malloc(sizeof(double) * (size_t)(*nz > 0 ? *nz : 1));
  for (int i = 0; i < *nz; i++) {
    int a = 0, b = 0, iv = 0; double x = 0, xi = 0;
    if (k->real) { if (fscanf(f, "%d %d %lg\n", &a, &b, &x) != 3) break; }
    else if (k->cplx) { if (fscanf(f, "%d %d %lg %lg\n", &a, &b, &x, &xi) != 4) break; }
    else if (k->integer) { if (fscanf(f, "%d %d %d\n", &a, &b, &iv) != 3) break; x = iv; }
    else { if (fscanf(f, "%d %d\n", &a, &b) != 2) break; x = 1.0; }
    (*ri)[i] = a - 1; (*cj)[i] = b - 1; (*vv)[i] = x;
  }
  fclose(f);
  return 0;
}
ORACLE_API int oracle_mtx_info(const char* fn, int* m, int* n, int* nnz, int* sym) {
  mm_kind k; int nz, *ri, *cj; double* vv;
  int rc = mm_load(fn, &k, m, n, &nz, &ri, &cj, &vv);
  if (rc) return rc;
  long tot = nz;
  if (k.sym) for (int i = 0; i < nz; i++) if (ri[i] != cj[i]) tot++;
  *nnz = (int)tot; *sym = k.sym;
  free(ri); free(cj); free(vv);
  return 0;
}
ORACLE_API int oracle_mtx_data(const char* fn, int* rp, int* ci, double* v) {
  mm_kind k; int m, n, nz, *ri, *cj; double* vv;
  int rc = mm_load(fn, &k, &m, &n, &nz, &ri, &cj, &vv);
  if (rc) return rc;
  int* cnt = (int*)calloc((size_t)m + 1, sizeof(int));
  for (int i = 0; i < nz; i++) cnt[ri[i]]++;
  if (k.sym) for (int i = 0; i < nz; i++) if (ri[i] != cj[i]) cnt[cj[i]]++;
  int acc = 0;
  for (int i = 0; i <= m; i++) { int c = cnt[i]; rp[i] = acc; acc += c; }
  memset(cnt, 0, sizeof(int) * ((size_t)m + 1));
  for (int i = 0; i < nz; i++) {
    int r = ri[i], c = cj[i];
    int o = rp[r] + cnt[r]++; ci[o] = c; v[o] = vv[i];
    if (k.sym && r != c) { o = rp[c] + cnt[c]++; ci[o] = r; v[o] = vv[i]; }
  }
  free(cnt); free(ri); free(cj); free(vv);
  return 0;
}

/* ---- format conversion --------------------------------------------------------
 * matrix.h:40-64 CsrToCsc: column histogram, exclusive scan (utility.h:276-288), then the entries
 * are dropped into their columns in ROW order, so a column keeps CSR order.  CscToCsr
 * (matrix.h:67-91) is the same loop with rows and columns exchanged: call this with (n, m, colptr,
 * rowidx, ...). */
#define ORACLE_CSR2CSC(SFX, T)                                                                          \
  ORACLE_API void oracle_csr2csc_##SFX(int m, int n, int nnz, const int* rp, const int* ci, const T* v, \
                                       int* cscRowIdx, int* cscColPtr, T* cscVal) {                     \
    memset(cscColPtr, 0, sizeof(int) * ((size_t)n + 1));                                                \
    for (int i = 0; i < nnz; i++) cscColPtr[ci[i]]++;                                                   \
    int run = 0;                                                                                        \
    for (int c = 0; c <= n; c++) { int t = cscColPtr[c]; cscColPtr[c] = run; run += t; }                \
    int* incr = (int*)malloc(sizeof(int) * ((size_t)n + 1));                                            \
    memcpy(incr, cscColPtr, sizeof(int) * ((size_t)n + 1));                                             \
    for (int row = 0; row < m; row++)                                                                   \
      for (int j = rp[row]; j < rp[row + 1]; j++) {                                                     \
        int col = ci[j];                                                                                \
        cscRowIdx[incr[col]] = row;                                                                     \
        cscVal[incr[col]] = v[j];                                                                       \
        incr[col]++;                                                                                    \
      }                                                                                                 \
    free(incr);                                                                                         \
  }
ORACLE_CSR2CSC(f64, double)
ORACLE_CSR2CSC(f32, float)

/* matrix.h:221-237 sortByRow: qsort of (row, col, val) records with cmp_func (matrix.h:100-115), which
 * orders by (row, col) and calls equal pairs equal — their relative order is whatever qsort does.
 * Restated as a STABLE merge sort by (row, col) (equal pairs keep input order), the one well-defined
 * choice; identical to the reference wherever (row, col) pairs are distinct. */
typedef struct { int r, c; double v; long pos; } oracle_coo_t;
static int oracle_coo_cmp(const void* a, const void* b) {
  const oracle_coo_t *x = (const oracle_coo_t*)a, *y = (const oracle_coo_t*)b;
  if (x->r != y->r) return x->r > y->r ? 1 : -1;
  if (x->c != y->c) return x->c > y->c ? 1 : -1;
  return x->pos > y->pos ? 1 : (x->pos < y->pos ? -1 : 0);
}
ORACLE_API void oracle_coo_sort_f64(long nnz, int* r, int* c, double* v) {
  oracle_coo_t* a = (oracle_coo_t*)malloc(sizeof(oracle_coo_t) * (size_t)(nnz > 0 ? nnz : 1));
  for (long i = 0; i < nnz; i++) { a[i].r = r[i]; a[i].c = c[i]; a[i].v = v[i]; a[i].pos = i; }
  qsort(a, (size_t)nnz, sizeof(oracle_coo_t), oracle_coo_cmp);
  for (long i = 0; i < nnz; i++) { r[i] = a[i].r; c[i] = a[i].c; v[i] = a[i].v; }
  free(a);
}

ORACLE_API int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
