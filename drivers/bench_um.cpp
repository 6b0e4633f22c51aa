// drivers/bench_um.cpp — the managed-memory variant (include/matrix_um.h, sblas_um.h) at real sizes: what the
// tuning hooks of matrix_um.h:529-614 (cudaMemAdvise + cudaMemPrefetchAsync) buy, as the reference's
// profiling.sh:3-21 asks of webbase-1M / s3dkq4m2 — here on the synthetic configs (no network for SuiteSparse).
//
//   ./bench_um spmv|spmm2 workload n_gpu tuning(0|1) [reps=3] [scale_down=1]
//     workload  c2 (Laplacian 4M rows) | b32 (banded 16M x 32) | c3v | b32m (banded 16M x 32, k = 64)
//
// Flow of spmv_test_um.cu:37-85: managed A / x / y built on the host, sync2gpu (aliases, nothing copied),
// tuning = 1: applyGpuTuning (advise + prefetch) before the op, removeGpuTuning / applyCpuTuning after it.
// Timed with cpu_timer (wall clock, as the reference): the hooks, the FIRST op call (tuning = 0: pages fault over
// PCIe inside the kernel), the following calls (resident), and the first host read of the result.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "harness.h"
#include "sblas_um.h"

namespace {
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
}  // namespace

int main(int argc, char* argv[]) {
  const drv::Args a{argc, argv};
  if (argc < 5) return a.usage("./bench_um spmv|spmm2 c2|b32|c3v|b32m n_gpu tuning [reps] [scale_down]");
  const std::string op = a.str(1), wl = a.str(2);
  const unsigned n_gpu = a.count(3);
  const bool tuning = a.flag(4);
  const int reps = argc > 5 ? (int)a.count(5) : 3;
  const int sd = argc > 6 ? std::max(1, (int)a.count(6)) : 1;
  const bool spmv = op == "spmv";
  const int k = spmv ? 1 : 64;
  int32_t rows, nx = 0, ny = 0;
  int64_t nnz;
  const bool lap = wl == "c2", uni = wl == "c3v" || wl == "c3";
  if (lap) { nx = 2048; ny = 2048 / sd; rows = nx * ny; nnz = sblas_synth_laplace2d_nnz(nx, ny); }
  else { rows = (1 << 24) / sd; nnz = (int64_t)rows * 32; }

  cpu_timer t;
  // ---- A generated on GPU 0 by the library's generators, then adopted into managed memory -------------------
  int32_t *d_rp = NULL, *d_ci = NULL;
  double* d_v = NULL;
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(int32_t) * ((size_t)rows + 1), (void**)&d_rp));
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(int32_t) * (size_t)nnz, (void**)&d_ci));
  SBLAS_SAFE_CALL(sblas_dev_alloc(0, sizeof(double) * (size_t)nnz, (void**)&d_v));
  if (lap) SBLAS_SAFE_CALL(sblas_synth_laplace2d(NULL, nx, ny, d_rp, d_ci, d_v, 1));
  else if (uni) SBLAS_SAFE_CALL(sblas_synth_uniform(NULL, rows, rows, 32, 211, d_rp, d_ci, d_v, 1));
  else SBLAS_SAFE_CALL(sblas_synth_banded(NULL, rows, 32, 211, d_rp, d_ci, d_v, 1));
  std::vector<int32_t> h_rp((size_t)rows + 1), h_ci((size_t)nnz);
  std::vector<double> h_v((size_t)nnz);
  SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_rp.data(), d_rp, sizeof(int32_t) * ((size_t)rows + 1)));
  SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_ci.data(), d_ci, sizeof(int32_t) * (size_t)nnz));
  SBLAS_SAFE_CALL(sblas_copy_d2h(0, h_v.data(), d_v, sizeof(double) * (size_t)nnz));
  sblas_dev_free(0, d_rp); sblas_dev_free(0, d_ci); sblas_dev_free(0, d_v);
  CsrSparseMatrix<int, double> A(rows, rows, h_rp.data(), h_ci.data(), h_v.data());
  printf("UM workload %s rows %d nnz %lld k %d op %s n_gpu %u tuning %d\n", wl.c_str(), rows, (long long)nnz, k, op.c_str(), n_gpu, (int)tuning);

  const int R = std::min<int>(rows, spmv ? 100000 : 1024);
  double t_tune = 0, t_first = 0, t_back = 0, t_read = 0, worst = 0;
  std::vector<double> t_next;
  bool ok = true;
  if (spmv) {
    DenseVector<int, double> x(rows, 1.0), y(rows, 1.0);
    fill_hash(x.val, (size_t)rows, 11);
    A.sync2gpu(n_gpu, segment);
    x.sync2gpu(n_gpu, replicate);
    y.sync2gpu(n_gpu, replicate);
    t.start_timer();
    if (tuning) { A.applyGpuTuning(); x.applyGpuTuning(); y.applyGpuTuning(false); }
    sblas_sync_gpus(n_gpu);
    t.stop_timer(); t_tune = t.measure();
    for (int r = 0; r <= reps; ++r) {
      t.start_timer();
      sblas_spmv_csr_v1<int, double>(&A, &x, &y, 1.0, r == 0 ? 1.0 : 0.0, n_gpu, tuning);
      sblas_sync_gpus(n_gpu);
      t.stop_timer();
      if (r == 0) t_first = t.measure(); else t_next.push_back(t.measure());
      if (r == 0) {   // validate the first result on the host (this read migrates / maps the pages of y back)
        cpu_timer tr; tr.start_timer();
        for (int i = 0; i < R; ++i) {
          double s = 0.0;
          for (int j = A.csrRowPtr[i]; j < A.csrRowPtr[i + 1]; ++j) s += A.csrVal[j] * x.val[A.csrColIdx[j]];
          const double ref = 1.0 + s;
          worst = std::max(worst, std::fabs(y.val[i] - ref) / (std::fabs(ref) + 1.0));
        }
        tr.stop_timer(); t_read = tr.measure();
        ok = worst <= 1e-12;
        if (tuning) { A.applyGpuTuning(); x.applyGpuTuning(); y.applyGpuTuning(false); sblas_sync_gpus(n_gpu); }  // the check pulled pages to the CPU
      }
    }
    t.start_timer();
    if (tuning) { A.removeGpuTuning(); x.removeGpuTuning(); y.removeGpuTuning(false); A.applyCpuTuning(); x.applyCpuTuning(); y.applyCpuTuning(); }
    y.sync2cpu(0, tuning);
    t.stop_timer(); t_back = t.measure();
  } else {
    DenseMatrix<int, double> B(rows, k, 0.0, col_major), C(rows, k, 1.0, col_major);
    fill_hash(B.val, (size_t)rows * k, 11);
    A.sync2gpu(n_gpu, segment);
    B.sync2gpu(n_gpu, replicate);
    C.sync2gpu(n_gpu, replicate);
    t.start_timer();
    if (tuning) { A.applyGpuTuning(); B.applyGpuTuning(); C.applyGpuTuning(false); }
    sblas_sync_gpus(n_gpu);
    t.stop_timer(); t_tune = t.measure();
    for (int r = 0; r <= reps; ++r) {
      t.start_timer();
      sblas_spmm_csr_v2<int, double>(&A, &B, &C, 1.0, r == 0 ? 1.0 : 0.0, n_gpu, tuning);
      sblas_sync_gpus(n_gpu);
      t.stop_timer();
      if (r == 0) t_first = t.measure(); else t_next.push_back(t.measure());
      if (r == 0) {
        cpu_timer tr; tr.start_timer();
        for (int n = 0; n < k; n += 7)
          for (int i = 0; i < R; ++i) {
            double s = 0.0;
            for (int j = A.csrRowPtr[i]; j < A.csrRowPtr[i + 1]; ++j) s += A.csrVal[j] * B.val[(size_t)n * rows + A.csrColIdx[j]];
            const double ref = 1.0 + s;
            worst = std::max(worst, std::fabs(C.val[(size_t)n * rows + i] - ref) / (std::fabs(ref) + 1.0));
          }
        tr.stop_timer(); t_read = tr.measure();
        ok = worst <= 1e-12;
        if (tuning) { A.applyGpuTuning(); B.applyGpuTuning(); C.applyGpuTuning(false); sblas_sync_gpus(n_gpu); }
      }
    }
    t.start_timer();
    if (tuning) { A.removeGpuTuning(); B.removeGpuTuning(); C.removeGpuTuning(false); A.applyCpuTuning(); B.applyCpuTuning(); C.applyCpuTuning(); }
    C.sync2cpu(0, tuning);
    t.stop_timer(); t_back = t.measure();
  }
  std::sort(t_next.begin(), t_next.end());
  const double med = t_next.empty() ? 0.0 : t_next[t_next.size() / 2];
  std::cout << "Validation = " << (ok ? "True" : "False") << "  (max |d|/(|ref|+1) = " << worst << ")" << std::endl;
  printf("UM_JSON {\"op\": \"%s\", \"workload\": \"%s\", \"n_gpu\": %u, \"tuning\": %d, \"rows\": %d, \"nnz\": %lld, \"k\": %d, \"hooks_before_ms\": %.3f, "
         "\"first_call_ms\": %.3f, \"later_calls_median_ms\": %.3f, \"host_check_after_first_call_ms\": %.3f, \"hooks_after_ms\": %.3f, \"validation\": %s}\n",
         op.c_str(), wl.c_str(), n_gpu, (int)tuning, rows, (long long)nnz, k, t_tune, t_first, med, t_read, t_back, ok ? "true" : "false");
  return ok ? 0 : 1;
}
