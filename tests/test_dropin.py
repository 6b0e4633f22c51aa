"""The C++ drop-in boundary: include/{sblas,spmv,spmm,matrix,matrix_um,kernel,utility,config}.h
over the C ABI, exercised through reference-shaped command-line drivers.

not gpu: the drivers build with plain g++ (no CUDA toolkit on the include path), every C-ABI
         symbol the headers use is exported, and without a GPU the drivers FAIL LOUDLY (there
         is no CPU fallback behind the ops).
gpu:     the drivers — ours (drivers/*.cpp) and the reference's OWN, unmodified spmv_test.cu /
         spmm_test.cu / *_um.cu compiled against our headers (oracle/_ref/refdrv_*, built here
         where /root/reference exists) — print "Validation = True" on ash85 for 1, 2, 4 GPUs.
"""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "drivers", "bin")
REFDRV = os.path.join(ROOT, "oracle", "_ref")
ASH = os.path.join(ROOT, "tests", "golden", "ash85.mtx")
LIB = os.path.join(ROOT, "s-blas_b200", "lib", "libsblas_b200.so")


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as ge
    ge._load_build().build_library()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "drivers")])
    return BIN


def _run(cmd, n_gpu_map=None, timeout=120):
    env = dict(os.environ)
    if n_gpu_map is not None:
        env["SBLAS_DEVICE_MAP"] = n_gpu_map
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_drivers_build_with_plain_gxx(built):
    for name in ("spmv_test", "spmm_test", "unit_test", "spmv_test_um", "spmm_test_um"):
        assert os.access(os.path.join(built, name), os.X_OK), name


def test_headers_only_use_exported_symbols(built):
    """Every sblas_* function named in the drop-in headers is exported by the library."""
    syms = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(sblas_[a-z0-9_]+)\b", syms))
    used = set()
    for h in ("utility.h", "matrix.h", "matrix_um.h", "spmv.h", "spmm.h", "kernel.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        used |= set(re.findall(r"\b(sblas_[a-z0-9_]+)\s*\(", src))
    local = {"sblas_safe_call_", "sblas_cuda_safe_call_", "sblas_cuda_check_error_", "sblas_sync_gpus",
             "sblas_default_dense_device_layout", "sblas_require_col_major_", "sblas_spmv_csr_cpu",
             "sblas_spmv_csr_v1", "sblas_spmm_csr_cpu", "sblas_spmm_csr_v1", "sblas_spmm_csr_v2", "sblas_range"}
    missing = sorted(u for u in used - local if u not in exported)
    assert not missing, missing


def test_usage_and_loud_failure_without_gpu(built):
    r = _run([os.path.join(built, "spmv_test")])
    assert r.returncode == 1 and "A_path alpha beta gpus" in r.stderr
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present: the loud-failure path is not reachable")
    r = _run([os.path.join(built, "spmv_test"), ASH, "1.0", "1.0", "2"])
    assert r.returncode != 0 and "sblas call failed" in r.stderr          # no silent CPU path
    assert "input matrix A: ( 85, 85 ) nnz = 523" in r.stdout            # the loader is host code


def test_reference_drivers_compile_unmodified(tmp_path):
    """/root/reference's own drivers against our headers (skipped where the checkout is absent)."""
    ref = "/root/reference"
    if not os.path.exists(os.path.join(ref, "spmv_test.cu")):
        pytest.skip("reference checkout not present")
    import shutil
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not present")
    for d in ("spmv_test", "spmm_test"):
        shutil.copy(os.path.join(ref, d + ".cu"), tmp_path / (d + ".cu"))
        r = subprocess.run([nvcc, "-ccbin", "/usr/bin/g++", "-O1", "-std=c++17", "-w", "-Wno-deprecated-gpu-targets",
                            "-I" + os.path.join(ROOT, "include"), "-c", "-o", str(tmp_path / (d + ".o")),
                            str(tmp_path / (d + ".cu"))], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]


# ------------------------------------------------------------------------------------- GPU ----
def _devmap():
    import torch
    n = max(torch.cuda.device_count(), 1)
    return ",".join(str(i % n) for i in range(16))


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpu", [1, 2, 4])
def test_spmv_driver(built, n_gpu):
    r = _run([os.path.join(built, "spmv_test"), ASH, "1.0", "1.0", str(n_gpu)], _devmap())
    assert r.returncode == 0 and "Validation = True" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpu", [1, 2, 4])
@pytest.mark.parametrize("method", [1, 2])
def test_spmm_driver(built, method, n_gpu):
    """README.md:76: ./spmm 1 ./ash85.mtx 64 1.0 1.0 4  (BASELINE config 1)."""
    r = _run([os.path.join(built, "spmm_test"), str(method), ASH, "64", "1.0", "1.0", str(n_gpu)], _devmap())
    assert r.returncode == 0 and "Validation = True" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["row", "col"])
def test_unit_test_driver(built, layout):
    """unit_test.cu:183-185: k=256, alpha=3, beta=4, 4 GPUs, plus the 1e-12 scaled bound."""
    env = dict(os.environ, SBLAS_DEVICE_MAP=_devmap(), SBLAS_DENSE_DEVICE_LAYOUT=layout)
    r = subprocess.run([os.path.join(built, "unit_test"), ASH, "4"], capture_output=True, text=True, timeout=300,
                       env=env, cwd=ROOT)
    assert r.returncode == 0 and "ALL PASSED" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("tuning", ["0", "1"])
@pytest.mark.parametrize("n_gpu", [1, 2])
def test_um_drivers(built, n_gpu, tuning):
    r = _run([os.path.join(built, "spmv_test_um"), ASH, "3.0", "4.0", str(n_gpu), tuning], _devmap())
    assert r.returncode == 0 and "Validation = True" in r.stdout, r.stdout + r.stderr
    for method in ("1", "2"):
        r = _run([os.path.join(built, "spmm_test_um"), method, ASH, "64", "1.0", "1.0", str(n_gpu), tuning], _devmap())
        assert r.returncode == 0 and "Validation = True" in r.stdout, (method, r.stdout + r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpu", [1, 2, 4])
def test_reference_own_drivers_on_our_headers(n_gpu):
    """The reference's unmodified spmv_test.cu / spmm_test.cu built against include/ (oracle/_ref/refdrv_*)."""
    exe = os.path.join(REFDRV, "refdrv_spmv_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/refdrv_* not built (needs /root/reference at build time)")
    r = _run([exe, ASH, "1.0", "1.0", str(n_gpu)], _devmap())
    assert "Validation = True" in r.stdout, r.stdout + r.stderr
    for method in ("1", "2"):
        r = _run([os.path.join(REFDRV, "refdrv_spmm_test"), method, ASH, "64", "1.0", "1.0", str(n_gpu)], _devmap())
        assert "Validation = True" in r.stdout, (method, r.stdout + r.stderr)
    r = _run([os.path.join(REFDRV, "refdrv_spmv_test_um"), ASH, "1.0", "1.0", str(n_gpu)], _devmap())
    assert "Validation = True" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_reference_own_unit_test_on_our_headers():
    """The reference's unmodified unit_test.cu (COO / CSR / CSC / dense containers, then SpMM method 1 and 2 and
    SpMV with k=256, alpha=3, beta=4 on 4 GPUs, unit_test.cu:176-187) built against include/."""
    exe = os.path.join(REFDRV, "refdrv_unit_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/refdrv_unit_test not built (needs /root/reference at build time)")
    env = dict(os.environ, SBLAS_DEVICE_MAP=_devmap())
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=env, cwd=os.path.dirname(ASH))
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("Validation = True") == 3 and "Validation = False" not in r.stdout, r.stdout + r.stderr
    assert "Building csc matrix from a csr matrix." in r.stdout
