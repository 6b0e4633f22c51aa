"""Builds s-blas_b200/lib/libsblas_b200.so for sm_100a, in-tree (so it travels to the GPU box).

    python s-blas_b200/build.py [--force]

nvcc cross-compiles without a GPU.  Objects are cached in s-blas_b200/_build/ keyed on the
source hashes, so a rebuild after touching one kernel file takes seconds.
"""
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libsblas_b200.so")

SOURCES = ["plan.cu", "spmv.cu", "spmm.cu", "spmm_rb.cu", "colpanel.cu", "dense.cu", "synth.cu", "mgpu.cu", "memory.cu", "convert.cu", "host.cpp"]
NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Wno-deprecated-gpu-targets", "-Wno-deprecated-declarations",
         "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wno-deprecated-declarations",
         "-I" + os.path.join(ROOT, "include"), "-I" + CSRC]


def _digest():
    h = hashlib.sha256(" ".join(FLAGS).encode())
    deps = sorted(os.listdir(CSRC)) + ["../../include/sblas_b200.h"]
    for f in deps:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p):
            with open(p, "rb") as fh:
                h.update(f.encode() + fh.read())
    return h.hexdigest()


def _compile(src):
    obj = os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o")
    cmd = [NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, " ".join(cmd), r.stderr[-4000:]))
    return obj


def build_library(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link libsblas_b200.so.  Returns its path."""
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = os.path.join(OBJ, "digest.txt")
    dg = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dg:
        return LIB
    if not os.path.exists(NVCC):
        if os.path.exists(LIB):  # GPU box without a rebuild need: use the shipped library
            return LIB
        raise RuntimeError("nvcc not found and no prebuilt library")
    with cf.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(_compile, SOURCES))
    cmd = [NVCC, "-shared", "-cudart", "static", "-o", LIB] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr[-4000:])
    with open(stamp, "w") as f:
        f.write(dg)
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
