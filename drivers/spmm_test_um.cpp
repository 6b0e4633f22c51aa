// drivers/spmm_test_um.cpp — managed-memory SpMM driver of the B200 build; command line of the
// reference's spmm_test_um.cu:   ./spmm_test_um method A_path B_width alpha beta gpus [tuning]
// Exit code 1 when validation fails.
#include <cstdlib>

#include "sblas_um.h"

static bool run(int method, const char* A_path, int b_width, double alpha, double beta, unsigned n_gpu, bool tuning) {
  cpu_timer load_timer, run_timer, run_cpu_timer;
  load_timer.start_timer();
  CsrSparseMatrix<int, double> A(A_path);
  DenseMatrix<int, double> B(A.width, b_width, col_major);
  DenseMatrix<int, double> C(A.height, b_width, 1, col_major);
  DenseMatrix<int, double> C_cpu(A.height, b_width, 1, col_major);
  if (method == 1) {  // spmm_test_um.cu: A replicated, B/C column slices alias the host arrays
    A.sync2gpu(n_gpu, replicate);
    B.sync2gpu(n_gpu, segment);
    C.sync2gpu(n_gpu, segment);
  } else {
    A.sync2gpu(n_gpu, segment);
    B.sync2gpu(n_gpu, replicate);
    C.sync2gpu(n_gpu, replicate);
  }
  if (tuning) {
    A.applyGpuTuning();
    B.applyGpuTuning();
    C.applyGpuTuning(false);
  }
  sblas_sync_gpus(n_gpu);
  load_timer.stop_timer();

  run_timer.start_timer();
  if (method == 1) sblas_spmm_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu, tuning);
  else sblas_spmm_csr_v2<int, double>(&A, &B, &C, alpha, beta, n_gpu, tuning);
  sblas_sync_gpus(n_gpu);
  run_timer.stop_timer();

  if (tuning) {
    A.removeGpuTuning();
    B.removeGpuTuning();
    C.removeGpuTuning(false);
    A.applyCpuTuning();
    B.applyCpuTuning();
    C.applyCpuTuning();
    C_cpu.applyCpuTuning();
    B.sync2cpu(0, tuning);
  }
  run_cpu_timer.start_timer();
  sblas_spmm_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  run_cpu_timer.stop_timer();

  C.sync2cpu(0, tuning);
  bool correct = check_equal(C_cpu.val, C.val, C.get_mtx_num());
  if (method == 2)
    for (unsigned i = 1; i < n_gpu; i++) correct = correct && check_equal(C_cpu.val, C.val_gpu[i], C.get_mtx_num());
  cout << "Validation = " << (correct ? "True" : "False") << endl;
  cout << "Load Time: " << load_timer.measure() << "ms." << endl;
  cout << n_gpu << "-GPUs Run Time: " << run_timer.measure() << " ms." << endl;
  cout << "CPU Run Time: " << run_cpu_timer.measure() << " ms." << endl;
  return correct;
}

int main(int argc, char* argv[]) {
  if (argc < 7) {
    cerr << "./spmm_test_um method A_path B_width alpha beta gpus tuning (optional, 0 by default)" << endl;
    return 1;
  }
  const int method = atoi(argv[1]);
  if (method != 1 && method != 2) {
    cerr << "method must be 1 (split B by columns) or 2 (split A by nnz)" << endl;
    return 1;
  }
  const bool tuning = argc > 7 ? atoi(argv[7]) == 1 : false;
  return run(method, argv[2], atoi(argv[3]), atof(argv[4]), atof(argv[5]), (unsigned)atoi(argv[6]), tuning) ? 0 : 1;
}
