"""sblas_b200 — B200-native (sm_100a) multi-GPU CSR SpMV / SpMM behind the S-BLAS interface.

The directory is named `s-blas_b200` (not importable by name); load it with
`__graft_entry__.load_package()`, which registers it as the module `sblas_b200`.

Layout:
  csrc/        hand-written CUDA kernels + the C ABI (include/sblas_b200.h)
  _lib.py      ctypes binding of lib/libsblas_b200.so          (no fallback)
  kernels.py   single-device calls + host partitioners
  matrix.py    CsrSparseMatrix / DenseMatrix / DenseVector      (matrix.h mirror)
  ops.py       sblas_spmv_csr_v1 / sblas_spmm_csr_v1 / _v2       (spmv.h / spmm.h mirror)
  dist.py      one-process-per-GPU variants over torch.distributed
  synth.py     device-side synthetic inputs of BASELINE.json's configs
"""
from . import _lib, dist, kernels, matrix, ops, synth, synth_np  # noqa: F401
from ._lib import (COL_MAJOR, FLAG_EXACT, FLAG_FMA, FLAG_HEAD_PARTIAL, FLAG_NO_L1, FLAG_NO_PREFETCH_B, FLAG_PREFETCH_B, FLAG_PREFETCH_C, FLAG_SPLIT_MERGE, FLAG_SPLIT_SEG, FLAG_SPLIT_SERIAL, FLAG_STITCH_BULK, FLAG_STITCH_DIRECT, FLAG_TAIL_PARTIAL,  # noqa: F401
                   ROW_MAJOR, SPMM_COLMAJOR, SPMM_ROWBLOCK, SPMM_ROWBLOCK_PIPE, SPMM_ROWBLOCK_TILEB, SPMM_ROWMAJOR, SPMV_COLPANEL, SPMV_PIPE, SPMV_SPLIT, SPMV_TILE, SPMV_VECTOR,
                   SblasError, make_opts)
from .matrix import (CooSparseMatrix, CscSparseMatrix, CscToCsr, CsrSparseMatrix, CsrToCsc, DenseMatrix,  # noqa: F401
                     DenseVector, col_major, none, replicate, row_major, segment)
from .ops import sblas_spmm_csr_v1, sblas_spmm_csr_v2, sblas_spmv_csr_v1  # noqa: F401

__version__ = "0.1.0"
