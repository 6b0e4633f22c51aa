// drivers/spmv_test_um.cpp — managed-memory SpMV driver of the B200 build; command line of the
// reference's spmv_test_um.cu:94-110:   ./spmv_test_um A_path alpha beta gpus [tuning]
// tuning = 1 applies the cudaMemAdvise / prefetch hints before the op and reverses them after
// it (spmv_test_um.cu:37-68).  Exit code 1 when validation fails.
#include <cstdlib>

#include "sblas_um.h"

static bool spmvCsrTest(const char* A_path, double alpha, double beta, unsigned n_gpu, bool tuning) {
  cpu_timer load_timer, run_timer, run_cpu_timer;
  load_timer.start_timer();
  CsrSparseMatrix<int, double> A(A_path);
  DenseVector<int, double> B(A.width, 1.);
  DenseVector<int, double> C(A.height, 1.);
  DenseVector<int, double> C_cpu(A.height, 1.);
  A.sync2gpu(n_gpu, segment);  // aliases into the managed arrays; only row pointers are new
  B.sync2gpu(n_gpu, replicate);
  C.sync2gpu(n_gpu, replicate);
  if (tuning) {
    A.applyGpuTuning();  // read-mostly + prefetch each GPU's range
    B.applyGpuTuning();
    C.applyGpuTuning(false);  // output: preferred location = its GPU
  }
  sblas_sync_gpus(n_gpu);
  load_timer.stop_timer();

  run_timer.start_timer();
  sblas_spmv_csr_v1<int, double>(&A, &B, &C, alpha, beta, n_gpu, tuning);
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
  sblas_spmv_csr_cpu<int, double>(&A, &B, &C_cpu, alpha, beta);
  run_cpu_timer.stop_timer();

  C.sync2cpu(0, tuning);  // replica 0 is C.val itself: nothing to copy
  bool correct = check_equal(C_cpu.val, C.val, C.get_vec_length());
  for (unsigned i = 1; i < n_gpu; i++) correct = correct && check_equal(C_cpu.val, C.val_gpu[i], C.get_vec_length());
  cout << "Validation = " << (correct ? "True" : "False") << endl;
  cout << "Load Time: " << load_timer.measure() << "ms." << endl;
  cout << n_gpu << "-GPUs Run Time: " << run_timer.measure() << " ms." << endl;
  cout << "CPU Run Time: " << run_cpu_timer.measure() << " ms." << endl;
  return correct;
}

int main(int argc, char* argv[]) {
  if (argc < 5) {
    cerr << "./spmv_test_um A_path alpha beta gpus tuning (optional, 0 by default)" << endl;
    return 1;
  }
  const bool tuning = argc > 5 ? atoi(argv[5]) == 1 : false;
  return spmvCsrTest(argv[1], atof(argv[2]), atof(argv[3]), (unsigned)atoi(argv[4]), tuning) ? 0 : 1;
}
