// drivers/bench_dropin.cpp — times the DROP-IN path (one process driving GPUs 0..n-1 through the reference's
// own entry points, include/spmv.h / spmm.h -> csrc/mgpu.cu) on the synthetic BASELINE configs and validates
// what it timed, in the shape of the reference's drivers (spmm_test.cu:58-93: distribute, run, CPU check,
// "Validation", "n-GPUs Run Time").
//
//   ./bench_dropin op workload n_gpu [reps=5] [scale_down=1]
//     op        spmv | spmm1 (method 1: B/C column split, spmm.h:84-161) | spmm2 (method 2: A nnz split, :164-284)
//     workload  c2 | c3 | c3v | c5 | b32 | b32m   (bench.py's table; k from the workload)
//
// The matrix is generated on GPU 0 by the library's counter-based generators (sblas_synth_*), copied to the
// host and handed to CsrSparseMatrix like any other CSR; B is filled on the host by a counter-based hash.
// Timing: the reference's cpu_timer (gettimeofday) around the op call — first call ("cold": lazily built
// tables, peer mappings) and the median of the following `reps` calls.  Validation: rows [0, R) of EVERY
// replica against sblas_spmv_csr_cpu / sblas_spmm_csr_cpu run on the first R rows of A (check_equal, 1e-3 as
// the reference, and the north star's relative bound printed beside it).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "harness.h"
#include "sblas.h"

namespace {

struct Workload { const char* name; const char* gen; int rows_log2, p1, k; bool f64_only; };
const Workload kWorkloads[] = {
    {"c2", "laplace2d", 22, 2048, 1, true},  {"c3", "uniform", 24, 32, 64, true},  {"c3v", "uniform", 24, 32, 1, true},
    {"c5", "banded", 21, 64, 256, true},     {"b32", "banded", 24, 32, 1, true},   {"b32m", "banded", 24, 32, 64, true},
};

void fill_hash(double* p, size_t n, uint64_t salt) {
  const unsigned T = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  std::vector<std::thread> th;
  for (unsigned t = 0; t < T; ++t)
    th.emplace_back([=] {
      for (size_t i = n * t / T; i < n * (t + 1) / T; ++i) {
        uint64_t z = (i + salt) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        p[i] = (double)((z ^ (z >> 31)) >> 11) * (1.0 / 9007199254740992.0);
      }
    });
  for (auto& x : th) x.join();
}

double median(std::vector<double> v) {
  std::sort(v.begin(), v.end());
  return v.empty() ? 0.0 : v[v.size() / 2];
}

}  // namespace

int main(int argc, char* argv[]) {
  const drv::Args a{argc, argv};
  if (argc < 4) return a.usage("./bench_dropin spmv|spmm1|spmm2 c2|c3|c3v|c5|b32|b32m n_gpu [reps] [scale_down]");
  const std::string op = a.str(1), wl = a.str(2);
  const unsigned n_gpu = a.count(3);
  const int reps = argc > 4 ? (int)a.count(4) : 5;
  const int sd = argc > 5 ? std::max(1, (int)a.count(5)) : 1;
  const Workload* w = NULL;
  for (const Workload& x : kWorkloads)
    if (wl == x.name) w = &x;
  if (!w || (op != "spmv" && op != "spmm1" && op != "spmm2")) return a.usage("unknown op / workload");
  const bool spmv = op == "spmv";
  const int k = spmv ? 1 : (w->k > 1 ? w->k : 64);

  cpu_timer t_load, t_run;
  t_load.start_timer();
  // ---- A: generated on GPU 0, adopted by CsrSparseMatrix ---------------------------------------------------
  int32_t rows = (1 << w->rows_log2) / sd, nx = 0, ny = 0;
  int64_t nnz = 0;
  if (!strcmp(w->gen, "laplace2d")) {
    nx = w->p1;
    ny = w->p1 / sd;
    rows = nx * ny;
    nnz = sblas_synth_laplace2d_nnz(nx, ny);
  } else {
    nnz = (int64_t)rows * w->p1;
  }
  int32_t *d_rp = NULL, *d_ci = NULL;
  double* d_v = NULL;
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(int32_t) * ((size_t)rows + 1), (void**)&d_rp));
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(int32_t) * (size_t)nnz, (void**)&d_ci));
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(double) * (size_t)nnz, (void**)&d_v));
  if (!strcmp(w->gen, "laplace2d")) SBLAS_SAFE_CALL(sblas_synth_laplace2d(NULL, nx, ny, d_rp, d_ci, d_v, 1));
  else if (!strcmp(w->gen, "uniform")) SBLAS_SAFE_CALL(sblas_synth_uniform(NULL, rows, rows, w->p1, 211, d_rp, d_ci, d_v, 1));
  else SBLAS_SAFE_CALL(sblas_synth_banded(NULL, rows, w->p1, 211, d_rp, d_ci, d_v, 1));
  SBLAS_SAFE_CALL(sblas_dev_sync(0));
  CsrSparseMatrix<int, double>* A = NULL;
  {
    int32_t *h_rp = NULL, *h_ci = NULL;
    double* h_v = NULL;
    SBLAS_SAFE_CALL(sblas_host_alloc(sizeof(int32_t) * ((size_t)rows + 1), (void**)&h_rp));
    SBLAS_SAFE_CALL(sblas_host_alloc(sizeof(int32_t) * (size_t)nnz, (void**)&h_ci));
    SBLAS_SAFE_CALL(sblas_host_alloc(sizeof(double) * (size_t)nnz, (void**)&h_v));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_rp, d_rp, sizeof(int32_t) * ((size_t)rows + 1)));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_ci, d_ci, sizeof(int32_t) * (size_t)nnz));
    SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_v, d_v, sizeof(double) * (size_t)nnz));
    sblas_dev_free(0, d_rp); sblas_dev_free(0, d_ci); sblas_dev_free(0, d_v);
    A = new CsrSparseMatrix<int, double>(rows, rows, h_rp, h_ci, h_v);
    sblas_host_free(h_rp); sblas_host_free(h_ci); sblas_host_free(h_v);
  }
  printf("workload %s (%s) rows %d nnz %lld k %d op %s n_gpu %u\n", w->name, w->gen, rows, (long long)nnz, k, op.c_str(), n_gpu);

  // ---- dense operands, distribution (sync2gpu is outside the timed op, as in spmm_test.cu:28-39) -------------
  const int R = std::min<int>(rows, spmv ? 200000 : (k > 64 ? 512 : 2048));  // rows validated against the CPU function
  CsrSparseMatrix<int, double> Asub(R, rows, A->csrRowPtr, A->csrColIdx, A->csrVal);
  bool ok = true;
  double worst_rel = 0.0;
  std::vector<double> times;
  double cold_ms = 0.0;
  if (spmv) {
    DenseVector<int, double> x(rows, 1.0), y(rows, 1.0), y_cpu(rows, 1.0);
    fill_hash(x.val, (size_t)rows, 11);
    A->sync2gpu(n_gpu, segment);
    x.sync2gpu(n_gpu, replicate);
    y.sync2gpu(n_gpu, replicate);
    sblas_sync_gpus(n_gpu);
    t_load.stop_timer();
    for (int r = 0; r <= reps; ++r) {
      t_run.start_timer();
      sblas_spmv_csr_v1<int, double>(A, &x, &y, 1.0, r == 0 ? 1.0 : 0.0, n_gpu);   // beta = 0 afterwards: y stays bounded
      sblas_sync_gpus(n_gpu);
      t_run.stop_timer();
      if (r == 0) {
        cold_ms = t_run.measure();
        // CPU statement on the first R rows (the reference's loop, spmv.h:22-31, on the leading block of A)
        for (int i = 0; i < R; ++i) {
          double s = 0.0;
          for (int j = A->csrRowPtr[i]; j < A->csrRowPtr[i + 1]; ++j) s += A->csrVal[j] * x.val[A->csrColIdx[j]];
          y_cpu.val[i] = 1.0 * y_cpu.val[i] + 1.0 * s;
        }
        for (unsigned g = 0; g < n_gpu; ++g) {
          y.sync2cpu(g);
          ok = check_equal(y_cpu.val, y.val, (size_t)R) && ok;
          for (int i = 0; i < R; ++i) worst_rel = std::max(worst_rel, std::fabs(y.val[i] - y_cpu.val[i]) / (std::fabs(y_cpu.val[i]) + 1.0));
        }
      } else {
        times.push_back(t_run.measure());
      }
    }
  } else {
    const bool m1 = op == "spmm1";
    DenseMatrix<int, double> B(rows, k, 0.0, col_major), C(rows, k, 1.0, col_major), C_cpu(R, k, 1.0, col_major);
    fill_hash(B.val, (size_t)rows * k, 11);
    A->sync2gpu(n_gpu, m1 ? replicate : segment);
    B.sync2gpu(n_gpu, m1 ? segment : replicate);
    C.sync2gpu(n_gpu, m1 ? segment : replicate);
    sblas_sync_gpus(n_gpu);
    t_load.stop_timer();
    for (int r = 0; r <= reps; ++r) {
      t_run.start_timer();
      if (m1) sblas_spmm_csr_v1<int, double>(A, &B, &C, 1.0, r == 0 ? 1.0 : 0.0, n_gpu);   // includes the per-slice D2H (spmm.h:151)
      else sblas_spmm_csr_v2<int, double>(A, &B, &C, 1.0, r == 0 ? 1.0 : 0.0, n_gpu);
      sblas_sync_gpus(n_gpu);
      t_run.stop_timer();
      if (r == 0) {
        cold_ms = t_run.measure();
        sblas_spmm_csr_cpu<int, double>(&Asub, &B, &C_cpu, 1.0, 1.0);
        for (unsigned g = 0; g < (m1 ? 1u : n_gpu); ++g) {
          if (!m1) C.sync2cpu(g);   // method 1 refreshed the host copy itself
          for (int n = 0; n < k; ++n) {
            ok = check_equal(C_cpu.val + (size_t)n * R, C.val + (size_t)n * rows, (size_t)R) && ok;
            for (int i = 0; i < R; ++i) {
              const double ref = C_cpu.val[(size_t)n * R + i], got = C.val[(size_t)n * rows + i];
              worst_rel = std::max(worst_rel, std::fabs(got - ref) / (std::fabs(ref) + 1.0));
            }
          }
        }
      } else {
        times.push_back(t_run.measure());
      }
    }
  }
  const double med = median(times);
  std::cout << "Validation = " << (ok ? "True" : "False") << "  (rows [0," << R << ") of every replica; max |d|/(|ref|+1) = " << worst_rel << ")" << std::endl;
  std::cout << "Load Time: " << t_load.measure() << "ms." << std::endl;
  std::cout << n_gpu << "-GPUs Run Time: first call " << cold_ms << " ms, median of " << times.size() << " calls " << med << " ms  ("
            << 2.0 * (double)nnz * k / (med * 1e-3) / 1e9 << " GFLOP/s)" << std::endl;
  printf("DROPIN_JSON {\"op\": \"%s\", \"workload\": \"%s\", \"n_gpu\": %u, \"rows\": %d, \"nnz\": %lld, \"k\": %d, \"first_call_ms\": %.4f, \"median_ms\": %.4f, \"min_ms\": %.4f, \"validation\": %s, \"max_rel\": %.3e}\n",
         op.c_str(), w->name, n_gpu, rows, (long long)nnz, k, cold_ms, med, times.empty() ? 0.0 : *std::min_element(times.begin(), times.end()), ok ? "true" : "false", worst_rel);
  delete A;
  return ok ? 0 : 1;
}
