// drivers/spmm_test.cpp — SpMM driver of the B200 build, same command line and output lines as
// the reference's spmm_test.cu:95-123:   ./spmm_test method A_path B_width alpha beta gpus
//   method 1: A replicated, B/C split by columns (spmm_test.cu:19-56)
//   method 2: A split by nnz, B/C replicated     (spmm_test.cu:58-93)
// Exit code 1 when validation fails.
#include <cstdlib>

#include "sblas.h"

static bool report(bool correct, cpu_timer& load, cpu_timer& run, cpu_timer& cpu, unsigned n_gpu) {
  cout << "Validation = " << (correct ? "True" : "False") << endl;
  cout << "Load Time: " << load.measure() << "ms." << endl;
  cout << n_gpu << "-GPUs Run Time: " << run.measure() << " ms." << endl;
  cout << "CPU Run Time: " << cpu.measure() << " ms." << endl;
  return correct;
}

static bool spmmCsrTest(const char* A_path, int b_width, double alpha, double beta, unsigned n_gpu) {
  cpu_timer load_timer, run_timer, run_cpu_timer;
  load_timer.start_timer();
  CsrSparseMatrix<int, double> A(A_path);
  DenseMatrix<int, double> B(A.width, b_width, col_major);  // rand(), seed 211
  DenseMatrix<int, double> C(A.height, b_width, 1, col_major);
  DenseMatrix<int, double> C_cpu(A.height, b_width, 1, col_major);
  A.sync2gpu(n_gpu, replicate);
  B.sync2gpu(n_gpu, segment);
  C.sync2gpu(n_gpu, segment);
  sblas_sync_gpus(n_gpu);
  load_timer.stop_timer();

  run_timer.start_timer();
  sblas_spmm_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu);  // host copy of C refreshed inside
  sblas_sync_gpus(n_gpu);
  run_timer.stop_timer();

  run_cpu_timer.start_timer();
  sblas_spmm_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  run_cpu_timer.stop_timer();
  return report(check_equal(C_cpu.val, C.val, C.get_mtx_num()), load_timer, run_timer, run_cpu_timer, n_gpu);
}

static bool spmmCsrTest2(const char* A_path, int b_width, double alpha, double beta, unsigned n_gpu) {
  cpu_timer load_timer, run_timer, run_cpu_timer;
  load_timer.start_timer();
  CsrSparseMatrix<int, double> A(A_path);
  DenseMatrix<int, double> B(A.width, b_width, col_major);
  DenseMatrix<int, double> C(A.height, b_width, 1, col_major);
  DenseMatrix<int, double> C_cpu(A.height, b_width, 1, col_major);
  A.sync2gpu(n_gpu, segment);
  B.sync2gpu(n_gpu, replicate);
  C.sync2gpu(n_gpu, replicate);
  sblas_sync_gpus(n_gpu);
  load_timer.stop_timer();

  run_timer.start_timer();
  sblas_spmm_csr_v2<int, double>(&A, &B, &C, alpha, beta, n_gpu);
  sblas_sync_gpus(n_gpu);
  run_timer.stop_timer();

  run_cpu_timer.start_timer();
  sblas_spmm_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  run_cpu_timer.stop_timer();

  bool correct = true;
  for (unsigned i = 0; i < n_gpu; i++) {  // every replica of C holds the result
    C.sync2cpu(i);
    correct = correct && check_equal(C_cpu.val, C.val, C.get_mtx_num());
  }
  return report(correct, load_timer, run_timer, run_cpu_timer, n_gpu);
}

int main(int argc, char* argv[]) {
  if (argc != 7) {
    cerr << "./spmm_test method A_path B_width alpha beta gpus" << endl;
    return 1;
  }
  const int method = atoi(argv[1]);
  const char* A_path = argv[2];
  const int b_width = atoi(argv[3]);
  const double alpha = atof(argv[4]), beta = atof(argv[5]);
  const unsigned gpus = (unsigned)atoi(argv[6]);
  bool ok;
  if (method == 1) ok = spmmCsrTest(A_path, b_width, alpha, beta, gpus);
  else if (method == 2) ok = spmmCsrTest2(A_path, b_width, alpha, beta, gpus);
  else {
    cerr << "method must be 1 (split B by columns) or 2 (split A by nnz)" << endl;
    return 1;
  }
  return ok ? 0 : 1;
}
