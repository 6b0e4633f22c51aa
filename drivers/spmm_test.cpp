// drivers/spmm_test.cpp — SpMM driver of the B200 build; command line and report lines of the
// reference's spmm_test.cu:95-123:   ./spmm_test method A_path B_width alpha beta gpus
//   method 1: A replicated, B and C split by columns, no exchange   (spmm.h:84-161)
//   method 2: A split by nnz, B and C replicated, stitched output   (spmm.h:164-284)
// C = beta*C + alpha*A*B with B random (seed 211) and C = 1, against the single-thread CPU statement.
// Exit code 1 when validation fails.
#include "harness.h"
#include "sblas.h"

namespace {
typedef drv::Phases<cpu_timer> Clock;

struct SpmmCase {
  const int method;
  CsrSparseMatrix<int, double> A;
  DenseMatrix<int, double> B, C, C_cpu;
  SpmmCase(int m, const char* path, int k)
      : method(m), A(path), B(A.width, k, col_major), C(A.height, k, 1, col_major), C_cpu(A.height, k, 1, col_major) {}

  void distribute(unsigned n_gpu) {
    const bool by_columns = method == 1;  // which operand is cut: the dense ones or the sparse one
    A.sync2gpu(n_gpu, by_columns ? replicate : segment);
    B.sync2gpu(n_gpu, by_columns ? segment : replicate);
    C.sync2gpu(n_gpu, by_columns ? segment : replicate);
    sblas_sync_gpus(n_gpu);
  }
  void multiply(double alpha, double beta, unsigned n_gpu) {
    if (method == 1) sblas_spmm_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu);  // refreshes the host C itself
    else sblas_spmm_csr_v2<int, double>(&A, &B, &C, alpha, beta, n_gpu);
    sblas_sync_gpus(n_gpu);
  }
  bool matches(unsigned n_gpu) {
    if (method == 1) return check_equal(C_cpu.val, C.val, C.get_mtx_num());
    bool ok = true;
    for (unsigned g = 0; g < n_gpu; ++g) {  // method 2: every replica of C holds the whole result
      C.sync2cpu(g);
      ok = check_equal(C_cpu.val, C.val, C.get_mtx_num()) && ok;
    }
    return ok;
  }
};
}  // namespace

int main(int argc, char* argv[]) {
  const drv::Args a{argc, argv};
  if (argc != 7) return a.usage("./spmm_test method A_path B_width alpha beta gpus");
  const int method = (int)a.count(1);
  if (method != 1 && method != 2) return a.usage("method must be 1 (split B by columns) or 2 (split A by nnz)");
  const double alpha = a.real(4), beta = a.real(5);
  const unsigned n_gpu = a.count(6);
  Clock clock;
  SpmmCase* c = NULL;
  clock.run(Clock::LOAD, [&] {
    c = new SpmmCase(method, a.str(2), (int)a.count(3));
    c->distribute(n_gpu);
  });
  clock.run(Clock::GPU, [&] { c->multiply(alpha, beta, n_gpu); });
  clock.run(Clock::CPU, [&] { sblas_spmm_csr_cpu<int, double>(&c->A, &c->B, &c->C_cpu, alpha, beta); });
  const bool ok = c->matches(n_gpu);
  clock.report(ok, n_gpu, /*gpu_line_first=*/true);
  delete c;
  return ok ? 0 : 1;
}
