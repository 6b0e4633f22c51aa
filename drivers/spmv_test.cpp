// drivers/spmv_test.cpp — SpMV driver of the B200 build; command line and report lines of the
// reference's spmv_test.cu:44-57:   ./spmv_test A_path alpha beta gpus
// y = beta*y + alpha*A*x with x = y = 1 on `gpus` GPUs (A split by nnz, x and y replicated) against
// the single-thread CPU statement on the same inputs.  Unlike the reference (whose mains always
// return 0) the exit code is 1 when validation fails.  Plain g++: everything below the headers is
// libsblas_b200.so.
#include "harness.h"
#include "sblas.h"

namespace {
typedef drv::Phases<cpu_timer> Clock;

struct SpmvCase {
  CsrSparseMatrix<int, double> A;
  DenseVector<int, double> x, y, y_cpu;
  explicit SpmvCase(const char* path) : A(path), x(A.width, 1.), y(A.height, 1.), y_cpu(A.height, 1.) {}

  void distribute(unsigned n_gpu) {  // spmv_test.cu:16-18
    A.sync2gpu(n_gpu, segment);      // nnz-balanced partition + upload + execution plans
    x.sync2gpu(n_gpu, replicate);
    y.sync2gpu(n_gpu, replicate);
    sblas_sync_gpus(n_gpu);
  }
  bool every_replica_matches(unsigned n_gpu) {  // matrix.h:714-726: all replicas are updated
    bool ok = true;
    for (unsigned g = 0; g < n_gpu; ++g) {
      y.sync2cpu(g);
      ok = check_equal(y_cpu.val, y.val, y.get_vec_length()) && ok;
    }
    return ok;
  }
};
}  // namespace

int main(int argc, char* argv[]) {
  const drv::Args a{argc, argv};
  if (argc != 5) return a.usage("./spmv_test A_path alpha beta gpus");
  const double alpha = a.real(2), beta = a.real(3);
  const unsigned n_gpu = a.count(4);
  Clock clock;
  SpmvCase* c = NULL;
  clock.run(Clock::LOAD, [&] {
    c = new SpmvCase(a.str(1));
    c->distribute(n_gpu);
  });
  clock.run(Clock::CPU, [&] { sblas_spmv_csr_cpu<int, double>(&c->A, &c->x, &c->y_cpu, alpha, beta); });
  clock.run(Clock::GPU, [&] {
    sblas_spmv_csr_v1<int, double>(&c->A, &c->x, &c->y, alpha, beta, n_gpu);
    sblas_sync_gpus(n_gpu);
  });
  const bool ok = c->every_replica_matches(n_gpu);
  clock.report(ok, n_gpu);
  delete c;
  return ok ? 0 : 1;
}
