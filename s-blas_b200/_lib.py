"""ctypes binding of libsblas_b200.so (the C ABI declared in include/sblas_b200.h).

There is NO fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libsblas_b200.so")

MAX_GPU = 16
# enums of include/sblas_b200.h
VARIANT_AUTO, SPMV_TILE, SPMV_VECTOR, SPMV_SPLIT, SPMV_PIPE, SPMV_COLPANEL = 0, 1, 2, 3, 4, 5
SPMM_ROWMAJOR, SPMM_COLMAJOR, SPMM_ROWBLOCK, SPMM_ROWBLOCK_PIPE, SPMM_ROWBLOCK_TILEB = 1, 2, 3, 4, 5
FLAG_FMA, FLAG_HEAD_PARTIAL, FLAG_TAIL_PARTIAL, FLAG_NO_L1 = 1, 2, 4, 8
FLAG_PREFETCH_B, FLAG_NO_PREFETCH_B = 16, 32
FLAG_SPLIT_MERGE = 64
FLAG_SPLIT_SERIAL = 128
FLAG_SPLIT_SEG = 256
FLAG_PREFETCH_C = 512
FLAG_EXACT = 1024
FLAG_STITCH_DIRECT = 2048
FLAG_STITCH_BULK = 4096
ROW_MAJOR, COL_MAJOR = 0, 1


class SblasError(RuntimeError):
    pass


class Opts(C.Structure):
    _fields_ = [("variant", C.c_int32), ("flags", C.c_int32), ("tile", C.c_int32),
                ("lanes", C.c_int32), ("n_replica", C.c_int32), ("stages", C.c_int32),
                ("replica", C.c_void_p * (MAX_GPU - 1)), ("carry", C.c_void_p)]


class CsrStats(C.Structure):
    _fields_ = [("nnz", C.c_int64), ("rows", C.c_int32), ("max_row_nnz", C.c_int32),
                ("empty_rows", C.c_int32), ("spmv_variant", C.c_int32), ("spmv_tile", C.c_int32),
                ("spmv_tiles", C.c_int32), ("spmv_smem_bytes", C.c_int32),
                ("spmm_tile", C.c_int32), ("spmm_tiles", C.c_int32)]


_lib = None


def lib():
    """Loads the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SblasError("%s not found: run `python s-blas_b200/build.py` (there is no CPU "
                             "fallback)" % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _lib.sblas_last_error.restype = C.c_char_p
        _lib.sblas_synth_laplace2d_nnz.restype = C.c_int64
        _lib.sblas_csr_find_row_i32.restype = C.c_int32
        _lib.sblas_launch_count.restype = C.c_ulonglong
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise SblasError("%s failed (%d): %s" % (what or "sblas call", rc,
                                                 lib().sblas_last_error().decode()))


def make_opts(variant=0, flags=0, tile=0, lanes=0, replicas=(), carry=None, stages=0):
    o = Opts()
    o.variant, o.flags, o.tile, o.lanes, o.stages = variant, flags, tile, lanes, stages
    o.n_replica = len(replicas)
    for i, p in enumerate(replicas):
        o.replica[i] = p
    o.carry = carry
    return o


def vp(t):
    """void* of a torch tensor / numpy array / int / None."""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return t.ctypes.data_as(C.c_void_p)


def stream_ptr(stream=None):
    """cudaStream_t of a torch stream (default: torch's current stream) as void*."""
    import torch
    s = torch.cuda.current_stream() if stream is None else stream
    return C.c_void_p(s.cuda_stream)
