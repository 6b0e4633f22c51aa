// drivers/spmv_test.cpp — SpMV driver of the B200 build, same command line and output lines as
// the reference's spmv_test.cu:44-57:   ./spmv_test A_path alpha beta gpus
// Builds A from a Matrix Market file, distributes it by nnz segments, runs the multi-GPU op
// and the CPU statement on the same inputs and prints "Validation = True/False".
// Unlike the reference (whose mains always return 0, unit_test.cu:186) the exit code is 1
// when validation fails.  Compiles with plain g++ (no CUDA toolkit needed): everything
// below the headers is libsblas_b200.so.
#include <cstdlib>

#include "sblas.h"

static bool spmvCsrTest(const char* A_path, double alpha, double beta, unsigned n_gpu) {
  cpu_timer load_timer, run_timer, run_cpu_timer;
  load_timer.start_timer();
  CsrSparseMatrix<int, double> A(A_path);
  DenseVector<int, double> B(A.width, 1.);
  DenseVector<int, double> C(A.height, 1.);
  DenseVector<int, double> C_cpu(A.height, 1.);
  A.sync2gpu(n_gpu, segment);  // nnz-balanced partition + upload + execution plans
  B.sync2gpu(n_gpu, replicate);
  C.sync2gpu(n_gpu, replicate);
  sblas_sync_gpus(n_gpu);
  load_timer.stop_timer();

  run_cpu_timer.start_timer();
  sblas_spmv_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  run_cpu_timer.stop_timer();

  run_timer.start_timer();
  sblas_spmv_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu);
  sblas_sync_gpus(n_gpu);
  run_timer.stop_timer();

  bool correct = true;
  for (unsigned i = 0; i < n_gpu; i++) {  // every replica holds the result (matrix.h:714-726)
    C.sync2cpu(i);
    correct = correct && check_equal(C_cpu.val, C.val, C.get_vec_length());
  }
  cout << "Validation = " << (correct ? "True" : "False") << endl;
  cout << "Load Time: " << load_timer.measure() << "ms." << endl;
  cout << "CPU Run Time: " << run_cpu_timer.measure() << " ms." << endl;
  cout << n_gpu << "-GPUs Run Time: " << run_timer.measure() << " ms." << endl;
  return correct;
}

int main(int argc, char* argv[]) {
  if (argc != 5) {
    cerr << "./spmv_test A_path alpha beta gpus" << endl;
    return 1;
  }
  return spmvCsrTest(argv[1], atof(argv[2]), atof(argv[3]), (unsigned)atoi(argv[4])) ? 0 : 1;
}
