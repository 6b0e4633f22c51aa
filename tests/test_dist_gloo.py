"""Host logic of the one-process-per-GPU row split (s-blas_b200/dist.py) on CPU: world_size 2 and
3 over gloo.  The CUDA kernel cannot run here, so the per-rank compute step is replaced by the
CPU oracle on the rank's local panel (oracle = checker standing in for the device, test only);
what is under test is the partition (panel_bounds / local_panel), the rebased local row
pointers, and the stitch (all-gather for equal panels, per-owner broadcasts for unequal ones)
reproducing the single-process result bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, case, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    import pyoracle
    pkg = ge.load_package()
    D = pkg.dist
    O = pyoracle.Oracle()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(case["seed"])          # same inputs on every rank
        h, w, k = case["h"], case["w"], case["k"]
        lens = rng.integers(0, case["maxlen"] + 1, h)
        if case.get("equal"):
            lens[:] = case["maxlen"]
        rp = np.zeros(h + 1, np.int32)
        rp[1:] = np.cumsum(lens)
        nnz = int(rp[-1])
        ci = rng.integers(0, w, nnz).astype(np.int32)
        v = rng.standard_normal(nnz)
        xB = rng.random(w * k)                              # row-major [w x k]
        y0 = rng.random(h * k)
        rb = D.panel_bounds(rp, world)
        assert rb[0] == 0 and rb[-1] == h and all(rb[i] <= rb[i + 1] for i in range(world))
        r0, r1, lrp, lci, lv = D.local_panel(rp, ci, v, rb, rank)
        assert lrp[0] == 0 and lrp[-1] == len(lci) == len(lv)

        def spmv_rows(rp_, ci_, v_, x_, y_, a, b):
            return O.spmv(np.ascontiguousarray(rp_, np.int32), ci_, v_, x_, y_, a, b)

        def spmm_rowmajor(rp_, ci_, v_, k_, B_rm, C_rm, a, b):
            hh = len(rp_) - 1
            Bc = np.ascontiguousarray(B_rm.reshape(w, k_).T).reshape(-1)
            Cc = np.ascontiguousarray(C_rm.reshape(hh, k_).T).reshape(-1)
            R = O.spmm(np.ascontiguousarray(rp_, np.int32), ci_, v_, k_, Bc, Cc, a, b, Bh=w, Ch=hh)
            return np.ascontiguousarray(np.asarray(R).reshape(k_, hh).T).reshape(-1)

        class CpuStandIn(D.RowSplitOp):
            """compute() on the oracle instead of the CUDA kernel; everything else inherited"""

            def compute(self, xB_t, out, alpha, beta):
                if self.r1 <= self.r0:
                    return
                lo, hi = self.r0 * self.ld, self.r1 * self.ld
                panel = out[lo:hi].numpy()
                if self.k == 1:
                    panel[:] = spmv_rows(self.rp, self.ci, self.val, xB_t.numpy(), panel.copy(), alpha, beta)
                else:
                    panel[:] = spmm_rowmajor(self.rp, self.ci, self.val, self.k, xB_t.numpy(), panel.copy(),
                                               alpha, beta)

        out = torch.from_numpy(y0.copy())
        op = CpuStandIn(h, w, rb, np.ascontiguousarray(lrp), np.ascontiguousarray(lci), np.ascontiguousarray(lv),
                        k=k, ld=k, stitch="nccl", plan=object())
        assert op.equal == (len({int(rb[i + 1] - rb[i]) for i in range(world)}) == 1)
        op(torch.from_numpy(xB), out, 1.5, -0.5)
        # single-process answer
        full = y0.copy()
        if k == 1:
            ref = spmv_rows(rp, ci, v, xB, full, 1.5, -0.5)
        else:
            ref = spmm_rowmajor(rp, ci, v, k, xB, full, 1.5, -0.5)
        ok = np.array_equal(out.numpy(), ref)
        open(os.path.join(out_dir, "rank%d.txt" % rank), "w").write("OK" if ok else "MISMATCH")
    finally:
        dist.destroy_process_group()


CASES = {
    "spmv_ragged_w2": dict(world=2, seed=1, h=501, w=300, k=1, maxlen=9),
    "spmv_equal_w2": dict(world=2, seed=2, h=512, w=300, k=1, maxlen=4, equal=True),
    "spmm_ragged_w2": dict(world=2, seed=3, h=257, w=100, k=8, maxlen=7),
    "spmm_ragged_w3": dict(world=3, seed=4, h=100, w=64, k=4, maxlen=5),
    "more_ranks_than_rows": dict(world=3, seed=5, h=2, w=10, k=1, maxlen=3),
}


@pytest.mark.parametrize("name", list(CASES))
def test_rowsplit_gloo(name, tmp_path):
    import torch.multiprocessing as mp
    case = CASES[name]
    world = case["world"]
    mp.spawn(_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / ("rank%d.txt" % r)).read() == "OK", (name, r)
