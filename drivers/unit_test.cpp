// drivers/unit_test.cpp — the reference's unit_test.cu cases for the SpMV/SpMM path, B200 build:
//   spmmCsrTest (256, 3.0, 4.0, 4)   unit_test.cu:66-104, :183   method 1
//   spmmCsrTest2(256, 3.0, 4.0, 4)   unit_test.cu:106-148, :184  method 2
//   spmvCsrTest (3.0, 4.0, 4)        unit_test.cu:150-175, :185
// on ./ash85.mtx (or argv[1]); argv[2] overrides the GPU count (default 4; with fewer physical
// devices set SBLAS_DEVICE_MAP, e.g. "0,0,0,0").  Besides the reference's 1e-3 absolute check,
// each case is also held to the north-star bound (1e-12 relative to |alpha|*|A||x| + |beta*y|).
// The COO/CSC container smoke tests of unit_test.cu:20-64 are outside the path.
#include <cmath>
#include <cstdlib>
#include <vector>

#include "sblas.h"

static const char* g_path = "./ash85.mtx";

// |alpha| * sum_j |a_ij| |b_jn| + |beta * c0|, col-major like C
static std::vector<double> bound_of(CsrSparseMatrix<int, double>& A, const double* B, int k, double alpha, double beta,
                                    double c0) {
  std::vector<double> bd((size_t)A.height * k);
  for (int n = 0; n < k; n++)
    for (int r = 0; r < A.height; r++) {
      double s = 0;
      for (int j = A.csrRowPtr[r]; j < A.csrRowPtr[r + 1]; j++)
        s += std::fabs(A.csrVal[j]) * std::fabs(B[(size_t)n * A.width + A.csrColIdx[j]]);
      bd[(size_t)n * A.height + r] = std::fabs(alpha) * s + std::fabs(beta * c0) + 1e-300;
    }
  return bd;
}

static bool spmmCase(int method, int b_width, double alpha, double beta, unsigned n_gpu) {
  CsrSparseMatrix<int, double> A(g_path);
  DenseMatrix<int, double> B(A.width, b_width, col_major);
  DenseMatrix<int, double> C(A.height, b_width, 1, col_major);
  DenseMatrix<int, double> C_cpu(A.height, b_width, 1, col_major);
  if (method == 1) {
    A.sync2gpu(n_gpu, replicate);
    B.sync2gpu(n_gpu, segment);
    C.sync2gpu(n_gpu, segment);
    sblas_spmm_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu);
  } else {
    A.sync2gpu(n_gpu, segment);
    B.sync2gpu(n_gpu, replicate);
    C.sync2gpu(n_gpu, replicate);
    sblas_spmm_csr_v2<int, double>(&A, &B, &C, alpha, beta, n_gpu);
  }
  sblas_sync_gpus(n_gpu);
  sblas_spmm_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  const std::vector<double> bd = bound_of(A, B.val, b_width, alpha, beta, 1.0);
  bool correct = true;
  for (unsigned i = 0; i < (method == 1 ? 1u : n_gpu); i++) {
    if (method == 2) C.sync2cpu(i);
    correct = correct && check_equal(C_cpu.val, C.val, C.get_mtx_num()) &&
              check_equal_scaled(C_cpu.val, C.val, bd.data(), C.get_mtx_num(), 1e-12);
  }
  cout << "spmm method " << method << ": Validation = " << (correct ? "True" : "False") << endl;
  return correct;
}

static bool spmvCase(double alpha, double beta, unsigned n_gpu) {
  CsrSparseMatrix<int, double> A(g_path);
  DenseVector<int, double> B(A.width);  // random, seed 211
  DenseVector<int, double> C(A.height, 1.);
  DenseVector<int, double> C_cpu(A.height, 1.);
  A.sync2gpu(n_gpu, segment);
  B.sync2gpu(n_gpu, replicate);
  C.sync2gpu(n_gpu, replicate);
  sblas_spmv_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu);
  sblas_sync_gpus(n_gpu);
  sblas_spmv_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  const std::vector<double> bd = bound_of(A, B.val, 1, alpha, beta, 1.0);
  bool correct = true;
  for (unsigned i = 0; i < n_gpu; i++) {
    C.sync2cpu(i);
    correct = correct && check_equal(C_cpu.val, C.val, C.get_vec_length()) &&
              check_equal_scaled(C_cpu.val, C.val, bd.data(), C.get_vec_length(), 1e-12);
  }
  cout << "spmv: Validation = " << (correct ? "True" : "False") << endl;
  return correct;
}

int main(int argc, char* argv[]) {
  if (argc > 1) g_path = argv[1];
  const unsigned n_gpu = argc > 2 ? (unsigned)atoi(argv[2]) : 4;
  bool ok = true;
  ok = spmmCase(1, 256, 3.0, 4.0, n_gpu) && ok;
  ok = spmmCase(2, 256, 3.0, 4.0, n_gpu) && ok;
  ok = spmvCase(3.0, 4.0, n_gpu) && ok;
  cout << (ok ? "ALL PASSED" : "FAILED") << endl;
  return ok ? 0 : 1;
}
