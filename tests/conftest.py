"""pytest configuration: registers the `gpu` marker and the shared fixtures.

`-m "not gpu"` : oracle vs golden vectors / reference, host logic, C-ABI symbol export.
`-m gpu`       : parity of the CUDA path (through the C ABI) against the oracle.
"""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import pyoracle
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def reference():
    import pyoracle
    if not pyoracle.Reference.available():
        pytest.skip("oracle/_ref/libsblas_ref.so not built (needs /root/reference)")
    return pyoracle.Reference()


@pytest.fixture(scope="session")
def golden():
    g = dict(np.load(os.path.join(GOLD, "golden.npz")))
    with open(os.path.join(GOLD, "golden.json")) as f:
        g["meta"] = json.load(f)
    g["ash85_path"] = os.path.join(GOLD, "ash85.mtx")
    return g


@pytest.fixture(scope="session")
def sblas():
    """The product package (hyphenated directory `s-blas_b200/`, imported as sblas_b200)."""
    import __graft_entry__ as ge
    try:
        import torch
        ndev = torch.cuda.device_count()
    except Exception:
        ndev = 0
    if ndev >= 1 and "SBLAS_DEVICE_MAP" not in os.environ:
        # fewer devices than the reference's tests use (4, unit_test.cu:183-185): fold the
        # logical GPUs onto the devices that exist (sblas_device_of)
        os.environ["SBLAS_DEVICE_MAP"] = ",".join(str(i % ndev) for i in range(16))
    return ge.load_package()


@pytest.fixture(scope="session")
def dev(sblas):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch.device("cuda:0")
