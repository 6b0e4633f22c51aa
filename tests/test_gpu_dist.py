"""One process per GPU (torchrun, NCCL): the row-split SpMV / SpMM of s-blas_b200/dist.py with the
peer-store, NVSwitch-multicast and NCCL stitches must leave the complete result, bit-identical to the
single-GPU op, on every rank (tools/check_dist.py).  Needs >= 2 GPUs; world-size 2/3 host logic is
covered on CPU by tests/test_dist_gloo.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_rowsplit_stitches_under_torchrun():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    n = min(n, 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "check_dist.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "check_dist OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
