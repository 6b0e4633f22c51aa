// drivers/unit_test.cpp — the reference's unit_test.cu cases for the SpMV/SpMM path, B200 build:
//   spmmCsrTest (256, 3.0, 4.0, 4)   unit_test.cu:66-104, :183   method 1
//   spmmCsrTest2(256, 3.0, 4.0, 4)   unit_test.cu:106-148, :184  method 2
//   spmvCsrTest (3.0, 4.0, 4)        unit_test.cu:150-175, :185
// on ./ash85.mtx (or argv[1]); argv[2] overrides the GPU count (default 4; with fewer physical
// devices set SBLAS_DEVICE_MAP, e.g. "0,0,0,0").  Besides the reference's 1e-3 absolute check,
// each case is also held to the north-star bound (1e-12 relative to |alpha|*|A||x| + |beta*y|).
// plus the container cases of unit_test.cu:20-51 (COO load + sort, CSR -> CSC, and back), each checked
// against plain CPU loops (the GPU conversions must be bit-identical to them).
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

// unit_test.cu:20-28: COO from file, sorted by (row, col); every (row, col, val) of the file must be there
static bool cooCase() {
  CooSparseMatrix<int, double> empty;
  CooSparseMatrix<int, double> coo(g_path, 2, segment);
  bool ok = coo.nnz > 0 && coo.n_gpu == 2 && coo.nnz_gpu[0] == coo.nnz / 2;
  for (int i = 1; i < coo.nnz && ok; i++)
    ok = coo.cooRowIdx[i - 1] < coo.cooRowIdx[i] ||
         (coo.cooRowIdx[i - 1] == coo.cooRowIdx[i] && coo.cooColIdx[i - 1] <= coo.cooColIdx[i]);
  std::vector<int> r((size_t)coo.nnz_gpu[1]);
  SBLAS_SAFE_CALL(sblas_copy_d2h(1, r.data(), coo.cooRowIdx_gpu[1], coo.get_gpu_nnz_idx_size(1)));
  for (int i = 0; i < coo.nnz_gpu[1] && ok; i++) ok = r[(size_t)i] == coo.cooRowIdx[coo.nnz / 2 + i];
  cout << "coo: Validation = " << (ok ? "True" : "False") << endl;
  return ok;
}

// unit_test.cu:42-51: CSC from CSR; checked against the serial loop of matrix.h:40-64, then converted back
static bool cscCase() {
  CscSparseMatrix<int, double> empty;
  CsrSparseMatrix<int, double> csr(g_path);
  CscSparseMatrix<int, double> csc(&csr);
  const int m = csr.height, n = csr.width, nnz = csr.nnz;
  std::vector<int> cp((size_t)n + 1, 0), ri((size_t)nnz), incr;
  std::vector<double> cv((size_t)nnz);
  for (int j = 0; j < nnz; j++) cp[(size_t)csr.csrColIdx[j] + 1]++;
  for (int c = 0; c < n; c++) cp[(size_t)c + 1] += cp[(size_t)c];
  incr = cp;
  for (int r = 0; r < m; r++)
    for (int j = csr.csrRowPtr[r]; j < csr.csrRowPtr[r + 1]; j++) {
      const int o = incr[(size_t)csr.csrColIdx[j]]++;
      ri[(size_t)o] = r;
      cv[(size_t)o] = csr.csrVal[j];
    }
  bool ok = csc.nnz == nnz && memcmp(cp.data(), csc.cscColPtr, sizeof(int) * ((size_t)n + 1)) == 0 &&
            memcmp(ri.data(), csc.cscRowIdx, sizeof(int) * (size_t)nnz) == 0 &&
            memcmp(cv.data(), csc.cscVal, sizeof(double) * (size_t)nnz) == 0;
  // and back: CSR with ascending columns inside every row, same multiset of entries per row
  std::vector<int> rp2((size_t)m + 1), ci2((size_t)nnz);
  std::vector<double> v2((size_t)nnz);
  CscToCsr<int, double>(n, m, nnz, csc.cscColPtr, csc.cscRowIdx, csc.cscVal, ci2.data(), rp2.data(), v2.data());
  ok = ok && memcmp(rp2.data(), csr.csrRowPtr, sizeof(int) * ((size_t)m + 1)) == 0;
  for (int r = 0; r < m && ok; r++)
    for (int j = rp2[(size_t)r] + 1; j < rp2[(size_t)r + 1] && ok; j++) ok = ci2[(size_t)j - 1] <= ci2[(size_t)j];
  cout << "csc: Validation = " << (ok ? "True" : "False") << endl;
  return ok;
}

int main(int argc, char* argv[]) {
  if (argc > 1) g_path = argv[1];
  const unsigned n_gpu = argc > 2 ? (unsigned)atoi(argv[2]) : 4;
  bool ok = true;
  ok = spmmCase(1, 256, 3.0, 4.0, n_gpu) && ok;
  ok = spmmCase(2, 256, 3.0, 4.0, n_gpu) && ok;
  ok = spmvCase(3.0, 4.0, n_gpu) && ok;
  ok = cooCase() && ok;
  ok = cscCase() && ok;
  cout << (ok ? "ALL PASSED" : "FAILED") << endl;
  return ok ? 0 : 1;
}
