// drivers/spmv_test_um.cpp — managed-memory SpMV driver of the B200 build; command line of the
// reference's spmv_test_um.cu:94-110:   ./spmv_test_um A_path alpha beta gpus [tuning]
// tuning = 1 applies the cudaMemAdvise / prefetch hints before the op and takes them back after it
// (spmv_test_um.cu:37-68).  Exit code 1 when validation fails.
#include "harness.h"
#include "sblas_um.h"

namespace {
typedef drv::Phases<cpu_timer> Clock;

struct SpmvUmCase {
  CsrSparseMatrix<int, double> A;
  DenseVector<int, double> x, y, y_cpu;
  explicit SpmvUmCase(const char* path) : A(path), x(A.width, 1.), y(A.height, 1.), y_cpu(A.height, 1.) {}

  void distribute(unsigned n_gpu, bool tuning) {
    A.sync2gpu(n_gpu, segment);  // aliases into the managed arrays; only the row pointers are new
    x.sync2gpu(n_gpu, replicate);
    y.sync2gpu(n_gpu, replicate);
    if (tuning) {
      A.applyGpuTuning();       // inputs: read-mostly + prefetch of each GPU's range
      x.applyGpuTuning();
      y.applyGpuTuning(false);  // output: preferred location = its GPU
    }
    sblas_sync_gpus(n_gpu);
  }
  void back_to_cpu(bool tuning) {  // spmv_test_um.cu:56-68
    if (!tuning) return;
    A.removeGpuTuning();
    x.removeGpuTuning();
    y.removeGpuTuning(false);
    A.applyCpuTuning();
    x.applyCpuTuning();
    y.applyCpuTuning();
    y_cpu.applyCpuTuning();
    x.sync2cpu(0, tuning);
  }
  bool every_replica_matches(unsigned n_gpu, bool tuning) {
    y.sync2cpu(0, tuning);  // replica 0 is y.val itself: nothing is copied
    bool ok = check_equal(y_cpu.val, y.val, y.get_vec_length());
    for (unsigned g = 1; g < n_gpu; ++g) ok = check_equal(y_cpu.val, y.val_gpu[g], y.get_vec_length()) && ok;
    return ok;
  }
};
}  // namespace

int main(int argc, char* argv[]) {
  const drv::Args a{argc, argv};
  if (argc < 5) return a.usage("./spmv_test_um A_path alpha beta gpus tuning (optional, 0 by default)");
  const double alpha = a.real(2), beta = a.real(3);
  const unsigned n_gpu = a.count(4);
  const bool tuning = a.flag(5);
  Clock clock;
  SpmvUmCase* c = NULL;
  clock.run(Clock::LOAD, [&] {
    c = new SpmvUmCase(a.str(1));
    c->distribute(n_gpu, tuning);
  });
  clock.run(Clock::GPU, [&] {
    sblas_spmv_csr_v1<int, double>(&c->A, &c->x, &c->y, alpha, beta, n_gpu, tuning);
    sblas_sync_gpus(n_gpu);
  });
  c->back_to_cpu(tuning);
  clock.run(Clock::CPU, [&] { sblas_spmv_csr_cpu<int, double>(&c->A, &c->x, &c->y_cpu, alpha, beta); });
  const bool ok = c->every_replica_matches(n_gpu, tuning);
  clock.report(ok, n_gpu, /*gpu_line_first=*/true);
  delete c;
  return ok ? 0 : 1;
}
