// host.cpp — integer/host layer of libsblas_b200.so: the reference-exact partitioners and the
// Matrix Market loader.  No CUDA in this file.
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

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
// NOT sorted within a row.
//
// SURVEY §8(f) f1 — the loader is the step right before the path (matrix.h:300-330) and the
// reference's per-entry fscanf dominates its "Load Time" on SuiteSparse-size inputs.  Here the file
// is read in one gulp and parsed by T host threads, order-preserving:
//   1. the entry region is cut into T byte ranges that begin at token starts; every thread counts
//      the whitespace-separated tokens that START in its range (the reference's fscanf is token-,
//      not line-based, so entries may be wrapped over lines in any way);
//   2. a prefix sum gives each range its first global token index g; entry e owns the tokens
//      [e*tpe, (e+1)*tpe) (tpe = 2 pattern, 3 real/integer, 4 complex), so a thread skips to its
//      first token with g % tpe == 0 and parses every entry that begins in its range (running
//      past the range end for the last one) straight into slot e of the coordinate arrays;
//   3. the CSR fill is a stable parallel counting sort: per-thread row histograms over contiguous
//      entry ranges, turned into per-thread start offsets, so every row keeps FILE order and a
//      mirrored entry sits right behind its original — bit-identical to the serial loop.
// SBLAS_MTX_THREADS (or sblas_mtx_set_threads) bounds T; T = 1 is the plain serial loop.
// sblas_mtx_info keeps the parsed file until the matching sblas_mtx_data call, so the usual
// info -> allocate -> data sequence (matrix.h:303-318) parses the file once, not twice.
namespace {
// Uninitialised array: std::vector would zero-fill (one extra pass and, worse, every page first touched by ONE
// thread); here the parser threads are the first to touch their own slots.
template <typename T>
class RawArray {
 public:
  RawArray() = default;
  RawArray(const RawArray&) = delete;
  RawArray& operator=(const RawArray&) = delete;
  RawArray(RawArray&& o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr; o.n_ = 0; }
  RawArray& operator=(RawArray&& o) noexcept {
    if (this != &o) { free(p_); p_ = o.p_; n_ = o.n_; o.p_ = nullptr; o.n_ = 0; }
    return *this;
  }
  ~RawArray() { free(p_); }
  bool resize(size_t n) {
    free(p_);
    p_ = n ? (T*)malloc(n * sizeof(T)) : nullptr;
    n_ = (n && !p_) ? 0 : n;
    return n == 0 || p_ != nullptr;
  }
  size_t size() const { return n_; }
  T* data() { return p_; }
  const T* data() const { return p_; }
  T& operator[](size_t i) { return p_[i]; }
  const T& operator[](size_t i) const { return p_[i]; }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};

struct Mtx {
  int m = 0, n = 0, sym = 0;
  RawArray<int32_t> ri, cj;
  RawArray<double> v;
};

int g_mtx_threads = 0;  // 0 = auto

static inline bool is_ws(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
static const char* skip_ws(const char* p, const char* e) {
  while (p < e && is_ws(*p)) ++p;
  return p;
}
static const char* skip_tok(const char* p, const char* e) {
  while (p < e && !is_ws(*p)) ++p;
  return p;
}
static const char* next_line(const char* p, const char* e) {
  while (p < e && *p != '\n') ++p;
  return p < e ? p + 1 : e;
}

static int mtx_thread_count(size_t bytes) {
  int t = g_mtx_threads;
  if (t <= 0) {
    if (const char* s = getenv("SBLAS_MTX_THREADS")) t = atoi(s);
  }
  if (t <= 0) {
    t = (int)std::thread::hardware_concurrency();
    if (t <= 0) t = 1;
    t = std::min(t, 64);
    t = (int)std::min<size_t>((size_t)t, std::max<size_t>(1, bytes >> 18));  // >= 256 KB of text per thread
  }
  return std::max(1, std::min(t, 256));
}

template <typename F>
static void run_threads(int T, F&& fn) {
  if (T <= 1) { fn(0); return; }
  std::vector<std::thread> th;
  th.reserve((size_t)T - 1);
  for (int t = 1; t < T; ++t) th.emplace_back([&fn, t] { fn(t); });
  fn(0);
  for (auto& x : th) x.join();
}

enum Field { F_REAL, F_CPLX, F_INT, F_PAT };

// Decimal index token -> long, exactly what strtol(p, &q, 10) returns for [+-]digits (the only form Matrix Market
// indices take); anything else (or > 18 digits) goes to strtol itself.  *end = first unparsed character.
static inline long fast_long(const char* p, const char** end) {
  const char* s = p;
  bool neg = false;
  if (*s == '-' || *s == '+') { neg = *s == '-'; ++s; }
  const char* d0 = s;
  unsigned long v = 0;
  while ((unsigned)(*s - '0') <= 9u) { v = v * 10u + (unsigned)(*s - '0'); ++s; }
  if (s == d0 || s - d0 > 18) {
    char* q = nullptr;
    const long r = strtol(p, &q, 10);
    *end = q > p ? q : p;
    return r;
  }
  *end = s;
  return neg ? -(long)v : (long)v;
}

// Decimal real token -> double.  Fast path (Clinger): a mantissa of at most 15 significant digits is exact in a
// double, so is 10^|e| for |e| <= 22, and ONE correctly rounded multiply or divide gives the correctly rounded
// result — bit-identical to strtod.  Everything else (longer mantissas such as %.17g output, large exponents,
// inf / nan / hex floats, malformed tokens) is handed to strtod unchanged.
static inline double fast_double(const char* p, const char** end) {
  static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                    1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  const char* s = p;
  bool neg = false;
  if (*s == '-' || *s == '+') { neg = *s == '-'; ++s; }
  unsigned long m = 0;
  int digits = 0, frac = 0;
  bool any = false;
  while ((unsigned)(*s - '0') <= 9u) {
    any = true;
    if (m || *s != '0') { m = m * 10u + (unsigned)(*s - '0'); ++digits; }
    ++s;
    if (digits > 15) goto slow;
  }
  if (*s == '.') {
    ++s;
    while ((unsigned)(*s - '0') <= 9u) {
      any = true;
      if (m || *s != '0') { m = m * 10u + (unsigned)(*s - '0'); ++digits; }
      ++frac;
      ++s;
      if (digits > 15 || frac > 300) goto slow;
    }
  }
  if (!any) goto slow;
  {
    int e10 = 0;
    if (*s == 'e' || *s == 'E') {
      const char* t = s + 1;
      bool eneg = false;
      if (*t == '-' || *t == '+') { eneg = *t == '-'; ++t; }
      if ((unsigned)(*t - '0') > 9u) goto slow;
      int ev = 0;
      while ((unsigned)(*t - '0') <= 9u) { ev = ev * 10 + (*t - '0'); ++t; if (ev > 1000) goto slow; }
      e10 = eneg ? -ev : ev;
      s = t;
    }
    if (!is_ws(*s) && *s != '\0') goto slow;  // trailing characters in the token (e.g. 'd' exponents): strtod decides
    e10 -= frac;
    if (e10 < -22 || e10 > 22) goto slow;
    double v = (double)m;  // exact: m < 10^15 < 2^53
    v = e10 < 0 ? v / kPow10[-e10] : v * kPow10[e10];
    *end = s;
    return neg ? -v : v;
  }
slow:
  char* q = nullptr;
  const double r = strtod(p, &q);
  *end = q > p ? q : p;
  return r;
}

// parse one entry whose first token starts at p; returns the position behind its last token
static inline const char* parse_entry(const char* p, const char* e, Field fld, int32_t* r, int32_t* c, double* x) {
  const char* q = p;
  long a = fast_long(p, &q); p = skip_tok(q, e); p = skip_ws(p, e);
  long b = fast_long(p, &q); p = skip_tok(q, e);
  double val = 1.0;
  if (fld == F_REAL) { p = skip_ws(p, e); val = fast_double(p, &q); p = skip_tok(q, e); }
  else if (fld == F_CPLX) {
    p = skip_ws(p, e); val = fast_double(p, &q); p = skip_tok(q, e);
    p = skip_ws(p, e); p = skip_tok(p, e);  // imaginary part dropped (mmio_highlevel.h:192)
  } else if (fld == F_INT) { p = skip_ws(p, e); val = (double)(int)fast_long(p, &q); p = skip_tok(q, e); }
  *r = (int32_t)(a - 1);
  *c = (int32_t)(b - 1);
  *x = val;
  return p;
}

static int load_mtx(const char* fn, Mtx& M) {
  const int fd = open(fn, O_RDONLY);
  if (fd < 0) { set_error("mtx: cannot open %s", fn); return -1; }
  struct stat fst;
  if (fstat(fd, &fst) != 0 || !S_ISREG(fst.st_mode)) { close(fd); set_error("mtx: cannot stat %s", fn); return -1; }
  const long sz = (long)fst.st_size;
  RawArray<char> buf;  // not zero-filled; one NUL behind the text: strtol/strtod stop there
  if (!buf.resize((size_t)std::max(sz, 0l) + 1)) { close(fd); set_error("mtx: out of memory reading %s", fn); return -1; }
  {  // the file is read by the same threads that parse it (pread of disjoint slices)
    const int TR = mtx_thread_count((size_t)std::max(sz, 0l));
    std::vector<int> bad((size_t)TR, 0);
    run_threads(TR, [&](int t) {
      size_t lo = (size_t)sz * (size_t)t / (size_t)TR, hi = (size_t)sz * (size_t)(t + 1) / (size_t)TR;
      while (lo < hi) {
        const ssize_t got = pread(fd, buf.data() + lo, hi - lo, (off_t)lo);
        if (got <= 0) { bad[(size_t)t] = 1; break; }
        lo += (size_t)got;
      }
    });
    close(fd);
    for (int t = 0; t < TR; ++t)
      if (bad[(size_t)t]) { set_error("mtx: short read"); return -1; }
  }
  buf[(size_t)std::max(sz, 0l)] = '\0';
  const char *p = buf.data(), *e = p + (size_t)std::max(sz, 0l);
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
  const Field fld = is_real ? F_REAL : is_cplx ? F_CPLX : is_int ? F_INT : F_PAT;
  const int tpe = is_pat ? 2 : is_cplx ? 4 : 3;  // tokens per entry
  M.sym = (!strcmp(w[4], "symmetric") || !strcmp(w[4], "hermitian")) ? 1 : 0;
  // size line, read the way mm_read_mtx_crd_size does (mmio.h:338-366): lines with '%' in column 0 are
  // comments; the first other line is tried as "M N nz" and consumed WHOLE (a fourth token on it is
  // dropped with the line, as fgets + sscanf drop it); if it is not three integers (a blank line, an
  // indented comment) it is discarded and the three integers are taken from the token stream behind it
  // (the reference's fscanf loop), the entries starting right behind the third one.
  while (p < e && *p == '%') p = next_line(p, e);
  if (p >= e) { set_error("mtx: premature end of file before the size line"); return -4; }
  long nz = 0;
  {
    const char* le = next_line(p, e);
    auto three = [&](const char* s, const char* lim, const char** after) -> bool {
      long vals[3];
      for (int i = 0; i < 3; ++i) {
        s = skip_ws(s, lim);
        if (s >= lim) return false;
        char* q = nullptr;
        vals[i] = strtol(s, &q, 10);
        if (q == s || q > lim) return false;
        s = q;
      }
      M.m = (int)vals[0]; M.n = (int)vals[1]; nz = vals[2];
      *after = s;
      return true;
    };
    const char* after = nullptr;
    if (three(p, le, &after)) {
      p = le;
    } else if (three(le, e, &after)) {
      p = after;
    } else {
      set_error("mtx: malformed size line");
      return -4;
    }
    if (M.m < 0 || M.n < 0 || nz < 0) { set_error("mtx: bad size line"); return -4; }
  }
  if (!M.ri.resize((size_t)nz) || !M.cj.resize((size_t)nz) || !M.v.resize((size_t)nz)) {
    set_error("mtx: out of memory for %ld entries", nz);
    return -1;
  }
  p = skip_ws(p, e);
  const size_t len = (size_t)(e - p);
  const int T = mtx_thread_count(len);
  // 1. range starts at token starts
  std::vector<const char*> start((size_t)T + 1);
  start[0] = p;
  start[(size_t)T] = e;
  for (int t = 1; t < T; ++t) {
    const char* q = p + len * (size_t)t / (size_t)T;
    if (q > p && !is_ws(q[-1])) q = skip_tok(q, e);  // inside a token: it belongs to the range before
    start[(size_t)t] = skip_ws(q, e);
  }
  for (int t = 1; t <= T; ++t) start[(size_t)t] = std::max(start[(size_t)t], start[(size_t)t - 1]);
  std::vector<long> ntok((size_t)T + 1, 0);
  run_threads(T, [&](int t) {
    const char *q = start[(size_t)t], *qe = start[(size_t)t + 1];
    long c = 0;
    while (q < qe) { ++c; q = skip_tok(q, e); q = skip_ws(q, e); }  // a token that starts here counts here
    ntok[(size_t)t + 1] = c;
  });
  // 2. first global token index of every range
  for (int t = 0; t < T; ++t) ntok[(size_t)t + 1] += ntok[(size_t)t];
  if (ntok[(size_t)T] < nz * tpe) {
    set_error("mtx: %ld entries announced, %ld found", nz, ntok[(size_t)T] / tpe);
    return -5;
  }
  run_threads(T, [&](int t) {
    const char *q = start[(size_t)t], *qe = start[(size_t)t + 1];
    long g = ntok[(size_t)t];
    while (q < qe && g % tpe != 0) { q = skip_tok(q, e); q = skip_ws(q, e); ++g; }  // tail of the previous range's entry
    long i = g / tpe;
    while (q < qe && i < nz) {
      q = parse_entry(q, e, fld, &M.ri[(size_t)i], &M.cj[(size_t)i], &M.v[(size_t)i]);
      q = skip_ws(q, e);
      ++i;
    }
  });
  return 0;
}

// the file parsed by the last sblas_mtx_info, kept for the sblas_mtx_data call that follows it
struct MtxCache {
  std::mutex mu;
  std::string key;
  Mtx M;
  bool valid = false;
} g_mtx_cache;

static std::string mtx_key(const char* fn) {
  struct stat st;
  if (stat(fn, &st) != 0) return std::string();
  char b[96];
  snprintf(b, sizeof b, "|%lld|%lld|%lld", (long long)st.st_size, (long long)st.st_mtime, (long long)st.st_ino);
  return std::string(fn) + b;
}
}  // namespace

extern "C" {

void sblas_mtx_set_threads(int n) { g_mtx_threads = n; }

int sblas_mtx_info(const char* fn, int32_t* m, int32_t* n, int32_t* nnz, int32_t* is_sym) {
  std::lock_guard<std::mutex> lk(g_mtx_cache.mu);
  Mtx& M = g_mtx_cache.M;
  g_mtx_cache.valid = false;
  M = Mtx();
  int rc = load_mtx(fn, M);
  if (rc) return rc;
  g_mtx_cache.key = mtx_key(fn);
  g_mtx_cache.valid = !g_mtx_cache.key.empty();
  int64_t tot = (int64_t)M.ri.size();
  if (M.sym) for (size_t i = 0; i < M.ri.size(); ++i) tot += (M.ri[i] != M.cj[i]);
  if (m) *m = M.m;
  if (n) *n = M.n;
  if (nnz) *nnz = (int32_t)tot;
  if (is_sym) *is_sym = M.sym;
  return 0;
}

int sblas_mtx_coo_info(const char* fn, int32_t* m, int32_t* n, int32_t* nz) {
  std::lock_guard<std::mutex> lk(g_mtx_cache.mu);
  Mtx& M = g_mtx_cache.M;
  g_mtx_cache.valid = false;
  M = Mtx();
  int rc = load_mtx(fn, M);
  if (rc) return rc;
  g_mtx_cache.key = mtx_key(fn);
  g_mtx_cache.valid = !g_mtx_cache.key.empty();
  if (m) *m = M.m;
  if (n) *n = M.n;
  if (nz) *nz = (int32_t)M.ri.size();
  return 0;
}

int sblas_mtx_coo_data(const char* fn, int32_t* rowidx, int32_t* colidx, double* val) {
  std::lock_guard<std::mutex> lk(g_mtx_cache.mu);
  Mtx local;
  Mtx* pM = &local;
  if (g_mtx_cache.valid && g_mtx_cache.key == mtx_key(fn)) {
    pM = &g_mtx_cache.M;
  } else {
    int rc = load_mtx(fn, local);
    if (rc) return rc;
  }
  const size_t nz = pM->ri.size();
  if (nz) {
    memcpy(rowidx, pM->ri.data(), nz * sizeof(int32_t));
    memcpy(colidx, pM->cj.data(), nz * sizeof(int32_t));
    memcpy(val, pM->v.data(), nz * sizeof(double));
  }
  g_mtx_cache.valid = false;
  g_mtx_cache.M = Mtx();
  return 0;
}

int sblas_mtx_data(const char* fn, int32_t* rowptr, int32_t* colidx, double* val) {
  std::lock_guard<std::mutex> lk(g_mtx_cache.mu);
  Mtx local;
  Mtx* pM = &local;
  if (g_mtx_cache.valid && g_mtx_cache.key == mtx_key(fn)) {
    pM = &g_mtx_cache.M;
  } else {
    int rc = load_mtx(fn, local);
    if (rc) return rc;
  }
  Mtx& M = *pM;
  struct Drop { ~Drop() { g_mtx_cache.valid = false; g_mtx_cache.M = Mtx(); } } drop;  // one data call per info call
  const size_t nz = M.ri.size();
  const size_t m1 = (size_t)M.m + 1;
  // threads for the counting sort: contiguous entry ranges, one row histogram each (<= 256 MB in all)
  int T = mtx_thread_count(nz * 24);
  T = (int)std::min<size_t>((size_t)T, std::max<size_t>(1, ((size_t)256 << 20) / (m1 * sizeof(int32_t))));
  std::vector<int32_t> hist((size_t)T * m1, 0);
  std::vector<int> bad((size_t)T, 0);
  std::vector<size_t> bad_at((size_t)T, 0);
  auto lo_of = [&](int t) { return nz * (size_t)t / (size_t)T; };
  run_threads(T, [&](int t) {
    int32_t* h = &hist[(size_t)t * m1];
    for (size_t i = lo_of(t); i < lo_of(t + 1); ++i) {
      const int32_t r = M.ri[i], c = M.cj[i];
      if (r < 0 || r >= M.m || c < 0 || c >= M.n) { if (!bad[(size_t)t]) { bad[(size_t)t] = 1; bad_at[(size_t)t] = i; } continue; }
      h[r]++;
      if (M.sym && r != c) {
        if (c >= M.m) { if (!bad[(size_t)t]) { bad[(size_t)t] = 2; bad_at[(size_t)t] = i; } continue; }
        h[c]++;
      }
    }
  });
  for (int t = 0; t < T; ++t) {
    if (bad[(size_t)t] == 1) { set_error("mtx: entry %zu out of range", bad_at[(size_t)t]); return -6; }
    if (bad[(size_t)t] == 2) { set_error("mtx: symmetric matrix is not square"); return -6; }
  }
  // row totals -> rowptr; hist[t][r] becomes thread t's first slot in row r
  int64_t acc = 0;
  for (size_t r = 0; r < (size_t)M.m; ++r) {
    rowptr[r] = (int32_t)acc;
    int64_t run = acc;
    for (int t = 0; t < T; ++t) {
      int32_t& h = hist[(size_t)t * m1 + r];
      const int32_t c = h;
      h = (int32_t)run;
      run += c;
    }
    acc = run;
  }
  rowptr[M.m] = (int32_t)acc;
  run_threads(T, [&](int t) {
    int32_t* h = &hist[(size_t)t * m1];
    for (size_t i = lo_of(t); i < lo_of(t + 1); ++i) {
      const int32_t r = M.ri[i], c = M.cj[i];
      int32_t o = h[r]++;
      colidx[o] = c; val[o] = M.v[i];
      if (M.sym && r != c) {
        o = h[c]++;
        colidx[o] = r; val[o] = M.v[i];
      }
    }
  });
  return 0;
}

}  // extern "C"
