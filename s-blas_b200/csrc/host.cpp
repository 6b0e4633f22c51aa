// host.cpp — integer/host layer of libsblas_b200.so: the reference-exact partitioners and the
// Matrix Market loader.  No CUDA in this file.
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "sblas_b200.h"

namespace sblas {
void set_error(const char* fmt, ...);
}
using sblas::set_error;

extern "C" {

// utility.h:293-300 — the row i with rowptr[i] <= idx < rowptr[i+1].  rowptr is
// non-decreasing, so that row is (upper_bound(rowptr, idx) - 1); it is automatically
// non-empty, which is what makes the reference's linear scan skip empty rows.
int32_t sblas_csr_find_row_i32(const int32_t* rowptr, int32_t height, int32_t idx) {
  if (!rowptr || height <= 0) return -1;
  const int32_t* ub = std::upper_bound(rowptr, rowptr + height + 1, idx);
  int64_t i = (ub - rowptr) - 1;
  if (i < 0 || i >= height) return -1;
  return (int32_t)i;
}

// matrix.h:360-369.  Arithmetic reproduced literally (including non-positive nnz_gpu and -1
// rows when there are more GPUs than nnz) except for avg_nnz, which is an integer ceil-div
// instead of `ceil((float)nnz / n_gpu)` (the float32 form loses nnz > 2^24).
int sblas_csr_partition_nnz_i32(const int32_t* rowptr, int32_t height, int32_t nnz, unsigned n_gpu,
                                int32_t* avg_out, int32_t* nnz_gpu, int32_t* start_row,
                                int32_t* stop_row) {
  if (!rowptr || !nnz_gpu || !start_row || !stop_row || n_gpu == 0 || height < 0 || nnz < 0) {
    set_error("partition_nnz: bad arguments");
    return 2;
  }
  const int64_t avg = ((int64_t)nnz + n_gpu - 1) / n_gpu;
  if (avg_out) *avg_out = (int32_t)avg;
  for (unsigned i = 0; i < n_gpu; i++) {
    const int64_t s = (int64_t)i * avg;
    const int64_t e = std::min<int64_t>((int64_t)(i + 1) * avg, nnz) - 1;
    nnz_gpu[i] = (int32_t)(e - s + 1);
    start_row[i] = sblas_csr_find_row_i32(rowptr, height, (int32_t)std::min<int64_t>(s, INT32_MAX));
    stop_row[i] = sblas_csr_find_row_i32(rowptr, height, (int32_t)e);
  }
  return 0;
}

// matrix.h:370-375
int sblas_csr_segment_rowptr_i32(const int32_t* rowptr, int32_t avg_nnz, unsigned i_gpu,
                                 int32_t start_row, int32_t stop_row, int32_t nnz_gpu_i,
                                 int32_t* out) {
  if (!rowptr || !out || start_row < 0 || stop_row < start_row) {
    set_error("segment_rowptr: bad arguments");
    return 2;
  }
  const int64_t num = (int64_t)stop_row - start_row + 2;
  const int32_t base = (int32_t)((int64_t)i_gpu * avg_nnz);
  out[0] = 0;
  for (int64_t k = 1; k < num - 1; k++) out[k] = rowptr[start_row + k] - base;
  out[num - 1] = nnz_gpu_i;
  return 0;
}

// matrix.h:557-567
int sblas_dense_split_i32(int32_t first_order, unsigned n_gpu, int32_t* avg_out, int32_t* dim_gpu) {
  if (!dim_gpu || n_gpu == 0 || first_order < 0) {
    set_error("dense_split: bad arguments");
    return 2;
  }
  const int64_t avg = ((int64_t)first_order + n_gpu - 1) / n_gpu;  // == ceil((double)fo / n)
  if (avg_out) *avg_out = (int32_t)avg;
  for (unsigned i = 0; i < n_gpu; i++)
    dim_gpu[i] = (int32_t)(std::min<int64_t>((int64_t)(i + 1) * avg, first_order) - (int64_t)i * avg);
  return 0;
}

// Row-aligned nnz-balanced panels: panel i starts at the first row whose first nnz index is
// >= i*avg_nnz, i.e. every row goes to the GPU on which the reference's split puts its start.
int sblas_csr_partition_rows_i32(const int32_t* rowptr, int32_t height, unsigned n_gpu,
                                 int32_t* row_begin) {
  if (!rowptr || !row_begin || n_gpu == 0 || height < 0) {
    set_error("partition_rows: bad arguments");
    return 2;
  }
  const int64_t nnz = rowptr[height];
  const int64_t avg = (nnz + n_gpu - 1) / n_gpu;
  row_begin[0] = 0;
  for (unsigned i = 1; i < n_gpu; i++) {
    const int64_t target = (int64_t)i * avg;
    if (target >= nnz) { row_begin[i] = height; continue; }
    row_begin[i] = (int32_t)(std::lower_bound(rowptr, rowptr + height, (int32_t)target) - rowptr);
  }
  row_begin[n_gpu] = height;
  return 0;
}

}  // extern "C"

// ---- Matrix Market (coordinate) -------------------------------------------------------------
// Semantics of mmio_highlevel.h:7-127 / :130-281: pattern -> 1.0, integer read as int, complex
// keeps the real part, symmetric or hermitian inputs mirrored (skew-symmetric is not), rows
// filled by a counting sort in file order (mirrored entry right behind its original), columns
// NOT sorted within a row.  The file is read in one gulp and tokenised by hand (the reference's
// per-entry fscanf dominates its "Load Time" on SuiteSparse-size inputs).
namespace {
struct Mtx {
  int m = 0, n = 0, sym = 0;
  std::vector<int32_t> ri, cj;
  std::vector<double> v;
};

static const char* skip_ws(const char* p, const char* e) {
  while (p < e && isspace((unsigned char)*p)) ++p;
  return p;
}
static const char* next_line(const char* p, const char* e) {
  while (p < e && *p != '\n') ++p;
  return p < e ? p + 1 : e;
}

static int load_mtx(const char* fn, Mtx& M) {
  FILE* f = fopen(fn, "rb");
  if (!f) { set_error("mtx: cannot open %s", fn); return -1; }
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::string buf((size_t)std::max(sz, 0l), '\0');
  if (sz > 0 && fread(&buf[0], 1, (size_t)sz, f) != (size_t)sz) { fclose(f); set_error("mtx: short read"); return -1; }
  fclose(f);
  const char *p = buf.data(), *e = p + buf.size();
  // banner: %%MatrixMarket matrix coordinate <type> <symmetry>
  char w[5][64] = {{0}};
  {
    const char* le = next_line(p, e);
    std::string line(p, le);
    if (sscanf(line.c_str(), "%63s %63s %63s %63s %63s", w[0], w[1], w[2], w[3], w[4]) != 5 ||
        strncmp(w[0], "%%MatrixMarket", 14) != 0) {
      set_error("mtx: could not process Matrix Market banner");
      return -2;
    }
    for (int i = 1; i < 5; ++i) for (char* q = w[i]; *q; ++q) *q = (char)tolower((unsigned char)*q);
    p = le;
  }
  if (strcmp(w[1], "matrix") != 0 || strcmp(w[2], "coordinate") != 0) { set_error("mtx: only 'matrix coordinate' is supported"); return -2; }
  const bool is_real = !strcmp(w[3], "real"), is_cplx = !strcmp(w[3], "complex"),
             is_int = !strcmp(w[3], "integer"), is_pat = !strcmp(w[3], "pattern");
  if (!(is_real || is_cplx || is_int || is_pat)) { set_error("mtx: unknown field type '%s'", w[3]); return -2; }
  M.sym = (!strcmp(w[4], "symmetric") || !strcmp(w[4], "hermitian")) ? 1 : 0;
  while (p < e && *p == '%') p = next_line(p, e);  // comments
  long nz = 0;
  {
    char* q = nullptr;
    M.m = (int)strtol(p, &q, 10); p = q;
    M.n = (int)strtol(p, &q, 10); p = q;
    nz = strtol(p, &q, 10); p = q;
    if (M.m < 0 || M.n < 0 || nz < 0) { set_error("mtx: bad size line"); return -4; }
  }
  M.ri.resize((size_t)nz); M.cj.resize((size_t)nz); M.v.resize((size_t)nz);
  for (long i = 0; i < nz; ++i) {
    char* q = nullptr;
    p = skip_ws(p, e);
    if (p >= e) { set_error("mtx: %ld entries announced, %ld found", nz, i); return -5; }
    long a = strtol(p, &q, 10); p = q;
    long b = strtol(p, &q, 10); p = q;
    double x = 1.0;
    if (is_real) { x = strtod(p, &q); p = q; }
    else if (is_cplx) { x = strtod(p, &q); p = q; (void)strtod(p, &q); p = q; }
    else if (is_int) { x = (double)(int)strtol(p, &q, 10); p = q; }
    M.ri[(size_t)i] = (int32_t)(a - 1);
    M.cj[(size_t)i] = (int32_t)(b - 1);
    M.v[(size_t)i] = x;
  }
  return 0;
}
}  // namespace

extern "C" {

int sblas_mtx_info(const char* fn, int32_t* m, int32_t* n, int32_t* nnz, int32_t* is_sym) {
  Mtx M;
  int rc = load_mtx(fn, M);
  if (rc) return rc;
  int64_t tot = (int64_t)M.ri.size();
  if (M.sym) for (size_t i = 0; i < M.ri.size(); ++i) tot += (M.ri[i] != M.cj[i]);
  if (m) *m = M.m;
  if (n) *n = M.n;
  if (nnz) *nnz = (int32_t)tot;
  if (is_sym) *is_sym = M.sym;
  return 0;
}

int sblas_mtx_data(const char* fn, int32_t* rowptr, int32_t* colidx, double* val) {
  Mtx M;
  int rc = load_mtx(fn, M);
  if (rc) return rc;
  const size_t nz = M.ri.size();
  std::vector<int32_t> cnt((size_t)M.m + 1, 0);
  for (size_t i = 0; i < nz; ++i) {
    if (M.ri[i] < 0 || M.ri[i] >= M.m || M.cj[i] < 0 || M.cj[i] >= M.n) { set_error("mtx: entry %zu out of range", i); return -6; }
    cnt[(size_t)M.ri[i]]++;
    if (M.sym && M.ri[i] != M.cj[i]) {
      if (M.cj[i] >= M.m) { set_error("mtx: symmetric matrix is not square"); return -6; }
      cnt[(size_t)M.cj[i]]++;
    }
  }
  int32_t acc = 0;
  for (int i = 0; i <= M.m; ++i) { int32_t c = cnt[(size_t)i]; rowptr[i] = acc; acc += c; }
  std::fill(cnt.begin(), cnt.end(), 0);
  for (size_t i = 0; i < nz; ++i) {
    const int32_t r = M.ri[i], c = M.cj[i];
    int32_t o = rowptr[r] + cnt[(size_t)r]++;
    colidx[o] = c; val[o] = M.v[i];
    if (M.sym && r != c) {
      o = rowptr[c] + cnt[(size_t)c]++;
      colidx[o] = r; val[o] = M.v[i];
    }
  }
  return 0;
}

}  // extern "C"
