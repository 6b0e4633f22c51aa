#!/usr/bin/env python
"""tools/check_dist.py — run under torchrun (one process per GPU): the row-split SpMV / SpMM of
s-blas_b200/dist.py with every stitch (peer stores over CUDA-IPC mappings, NVSwitch multicast
stores through torch symmetric memory, NCCL all-gather) against the same op computed whole on
this rank's GPU.  Every rank must end up with the complete, bit-identical y / C.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29533 tools/check_dist.py
Prints one line per case and `check_dist OK` on rank 0; exit code 1 on any mismatch."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("s-blas_b200")
K, D = pkg.kernels, pkg.dist


def banded(n, bw, dtype, seed):
    rng = np.random.default_rng(seed)
    c0 = np.clip(np.arange(n) - bw // 2, 0, n - bw)
    cols = c0[:, None] + np.arange(bw)[None, :]
    keep = rng.random(cols.shape) >= 0.1
    keep[:, 0] = True
    rp = np.zeros(n + 1, np.int32)
    rp[1:] = np.cumsum(keep.sum(1))
    ci = cols[keep].astype(np.int32)
    v = (rng.random(len(ci)) * 2 - 1).astype(dtype)
    return rp, ci, v


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    bad = 0
    for stitch in ("peer", "mc", "nccl"):
        for k in (1, 64):
            for steps in (1, 3):
                n = 20000 + 7 * world
                rp, ci, v = banded(n, 12, np.float64, 3)
                t = lambda a: torch.from_numpy(a).to(dev)   # noqa: E731
                drp, dci, dv = t(rp), t(ci), t(v)
                ld = k
                x = torch.from_numpy(np.random.default_rng(5).random(n * ld)).to(dev)
                # whole op on this GPU
                ref = torch.ones(n * ld, dtype=torch.float64, device=dev)
                plan = K.Plan(drp, len(ci))
                for _ in range(steps):
                    if k == 1:
                        K.spmv(drp, dci, dv, x, ref, 1.5, 0.5, plan=plan, cols=n)
                    else:
                        K.spmm(drp, dci, dv, k, pkg.ROW_MAJOR, x, ld, ref, ld, 1.5, 0.5, n, plan=plan)
                # row-split across the ranks
                rb = D.panel_bounds(rp, world)
                r0, r1, lrp, lci, lv = D.local_panel(drp, dci, dv, rb, rank)
                lci, lv = lci.clone(), lv.clone()
                peer_ptrs, mc_ptr, keep = None, 0, None
                mode = stitch
                if stitch == "mc":
                    keep = D.SymmBuffer(n * ld, torch.float64, dev)
                    out, peer_ptrs, mc_ptr = keep.tensor(), keep.peer_ptrs, keep.mc_ptr
                    if not mc_ptr:
                        mode = "peer"
                elif stitch == "peer":
                    keep = D.IpcBuffer(n * ld * 8, dev)
                    peer_ptrs = keep.open_peers()
                    out = keep.tensor(torch.float64, n * ld)
                else:
                    out = torch.empty(n * ld, dtype=torch.float64, device=dev)
                out.fill_(1.0)
                torch.cuda.synchronize()
                dist.barrier()
                op = D.RowSplitOp(n, n, rb, lrp.contiguous(), lci, lv, k=k, ld=ld if k > 1 else 1, stitch=mode,
                                  peer_ptrs=peer_ptrs, mc_ptr=mc_ptr)
                for _ in range(steps):
                    op(x, out, 1.5, 0.5)
                torch.cuda.synchronize()
                dist.barrier()
                torch.cuda.synchronize()
                ok = bool(torch.equal(out, ref))
                if getattr(op, "status", None) is not None and int(op.status.item()) != 0:
                    ok = False
                flag = torch.tensor([0 if ok else 1], device=dev)
                dist.all_reduce(flag)
                if rank == 0:
                    print("stitch=%s%s k=%d steps=%d world=%d : %s" % (
                        stitch, "" if mode == stitch else "(->%s: no multicast here)" % mode, k, steps, world,
                        "ok" if int(flag) == 0 else "MISMATCH on %d rank(s)" % int(flag)), flush=True)
                bad += int(flag)
                del op, out, keep
    if rank == 0:
        print("check_dist OK" if bad == 0 else "check_dist FAILED")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
