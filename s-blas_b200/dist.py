"""One-process-per-GPU execution of the A-row-split ops (SpMV v1 / SpMM method 2) over
torch.distributed — the plumbing bench.py uses at N > 1.

North star, method 2: "splits A into nnz-balanced row panels and stitches y/C with an NCCL
allgather over NVLink or with direct peer writes".  Rank r owns the row panel
[row_begin[r], row_begin[r+1]) (sblas_csr_partition_rows_i32: every row goes to the rank on
which the reference's nnz split, matrix.h:360-369, puts its first nnz; no row is shared, so
no reduction is needed).  x/B and y/C are replicated on every rank, as in the reference
(spmm_test.cu:68-70).  Two stitches:

  stitch="nccl"  kernel writes the local panel of the local replica; the panels are then
                 all-gathered in place (equal panels) or broadcast panel by panel;
  stitch="peer"  the replicas live in CUDA-IPC shared allocations; the kernel epilogue stores
                 each finished row into EVERY rank's replica over NVLink (sblasOpts.replica),
                 so there is no separate collective — only a device-side flag barrier
                 (sblas_peer_barrier) enqueued behind the kernel, no host synchronisation.
On CPU (gloo) the same host logic runs with a caller-supplied compute stand-in (tests only).
"""
import ctypes as C

import numpy as np

from . import kernels as K
from ._lib import ROW_MAJOR, check, lib, make_opts, stream_ptr


def _dist():
    import torch.distributed as dist
    return dist


def panel_bounds(rowptr_host, world):
    """row_begin[world+1] of the nnz-balanced row-aligned panels."""
    return K.csr_partition_rows(rowptr_host, world)


def local_panel(rowptr, colidx, val, row_begin, rank):
    """Slices rank's panel out of a full CSR (host numpy or device torch): local rowptr is
    rebased to start at 0 (matrix.h:372-375 without the shared-row special cases)."""
    r0, r1 = int(row_begin[rank]), int(row_begin[rank + 1])
    n0, n1 = int(rowptr[r0]), int(rowptr[r1])
    rp = rowptr[r0:r1 + 1] - n0
    return r0, r1, rp, colidx[n0:n1], val[n0:n1]


class IpcBuffer:
    """A device allocation every rank of the node can store into (cudaIpc*)."""

    def __init__(self, nbytes, device):
        import torch
        self.nbytes, self.device = int(nbytes), device
        self.ptr = C.c_void_p()
        self.handle = (C.c_ubyte * 64)()
        with torch.cuda.device(device):
            check(lib().sblas_ipc_alloc(C.c_size_t(self.nbytes), C.byref(self.ptr), self.handle), "ipc_alloc")
        self.peers = {}

    def open_peers(self):
        """Exchanges handles; afterwards self.peer_ptrs[r] is rank r's buffer mapped here."""
        import torch
        dist = _dist()
        rank, world = dist.get_rank(), dist.get_world_size()
        handles = [None] * world
        dist.all_gather_object(handles, bytes(self.handle))
        self.peer_ptrs = [None] * world
        with torch.cuda.device(self.device):
            for r in range(world):
                if r == rank:
                    self.peer_ptrs[r] = self.ptr.value
                    continue
                p = C.c_void_p()
                h = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                check(lib().sblas_ipc_open(h, C.byref(p)), "ipc_open")
                self.peer_ptrs[r] = p.value
        return self.peer_ptrs

    def tensor(self, dtype, numel):
        """torch view of the local buffer (no copy)."""
        import torch
        esz = torch.empty(0, dtype=dtype).element_size()

        class _Iface:
            pass
        o = _Iface()
        o.__cuda_array_interface__ = {"shape": (int(numel),), "typestr": "<f%d" % esz,
                                      "data": (self.ptr.value, False), "version": 3, "strides": None}
        return torch.as_tensor(o, device=self.device)

    def close(self):
        import torch
        with torch.cuda.device(self.device):
            for r, p in enumerate(getattr(self, "peer_ptrs", [])):
                if p is not None and p != self.ptr.value:
                    lib().sblas_ipc_close(C.c_void_p(p))
            if self.ptr:
                lib().sblas_ipc_free(self.ptr)
                self.ptr = C.c_void_p()


class SymmBuffer:
    """Output replica in torch symmetric memory (torch.distributed._symmetric_memory, CUDA backend):
    every rank maps every peer's buffer AND, on an NVSwitch box, one multicast address whose stores
    the switch replicates into all ranks' buffers (NVLS).  `mc_ptr` is 0 where the box has no
    multicast support; `peer_ptrs` always works (the same role as IpcBuffer.peer_ptrs)."""

    def __init__(self, numel, dtype, device):
        import torch.distributed._symmetric_memory as symm_mem
        dist = _dist()
        self.t = symm_mem.empty(int(numel), dtype=dtype, device=device)
        self.hdl = symm_mem.rendezvous(self.t, dist.group.WORLD)
        self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.mc_ptr = int(self.hdl.multicast_ptr) if getattr(self.hdl, "has_multicast_support", True) else 0

    def tensor(self):
        return self.t


def choose_stitch(row_begin, world):
    """The stitch for replicated y / C when the caller does not name one.  In-kernel peer stores are the
    fastest when the panels hold similar numbers of rows (every GPU sends (n-1) x its panel).  nnz-balanced
    panels of a skewed matrix do not: the GPU that owns the long sparse tail sends several times the mean,
    and its NVLink egress sets the step (R-MAT config 4, 8 GPUs: one panel holds 43 % of the rows, 0.312 ms).
    NVSwitch multicast stores send each panel ONCE and let the switch replicate it (0.172 ms)."""
    rows = [int(row_begin[i + 1]) - int(row_begin[i]) for i in range(world)]
    mean = max(sum(rows) / max(world, 1), 1.0)
    return "mc" if world >= 4 and max(rows) > 2.0 * mean else "peer"


class RowSplitOp:
    """y/C <- beta*y/C + alpha*A*x/B with A split into row panels across the ranks.

    rp/ci/val: this rank's local panel (device tensors).  `out` is the full replicated y
    (k=1) or row-major C [rows x ld] of this rank; with stitch="peer" it must be the tensor
    view of an IpcBuffer whose peers have been opened (pass peer_ptrs)."""

    def __init__(self, rows_total, cols, row_begin, rp, ci, val, k=1, ld=1, stitch="nccl",
                 peer_ptrs=None, plan=None, mc_ptr=0, flags=0):
        dist = _dist()
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rows_total, self.cols, self.row_begin = rows_total, cols, [int(x) for x in row_begin]
        self.rp, self.ci, self.val, self.k, self.ld = rp, ci, val, k, ld
        self.r0, self.r1 = self.row_begin[self.rank], self.row_begin[self.rank + 1]
        self.plan = plan if plan is not None else (K.Plan(rp, val.numel()) if self.r1 > self.r0 else None)
        if plan is None and self.plan is not None and not (k == 1 and ld == 1):
            self.plan.build_rowblocks(ci, val)      # SpMM: row-block mirror when it pays (sync2gpu work)
        elif plan is None and self.plan is not None:
            # SpMV: tile table + column statistic now (no call then synchronises), sliced-ELL mirror when it pays
            self.plan.prepare(ci, cols, val.element_size())
            self.plan.build_colpanels(ci, val, cols)
        self.stitch = stitch if self.world > 1 else "none"
        self.equal = len({self.row_begin[i + 1] - self.row_begin[i] for i in range(self.world)}) == 1
        self.opts = make_opts(flags=flags) if flags else None   # extra SBLAS_FLAG_* for the kernels (experiments)
        self.flags = None
        if self.stitch == "mc":
            # one store per finished row slice into the multicast address: the NVSwitch replicates it
            # into every rank's buffer (this rank's included), so a GPU sends |panel| bytes, not (n-1)x
            if not mc_ptr:
                raise RuntimeError("stitch='mc' needs a multicast address (SymmBuffer.mc_ptr is 0 on this box)")
            from ._lib import FLAG_STITCH_DIRECT   # multicast addresses take plain stores, not TMA bulk copies
            self.opts = make_opts(replicas=[mc_ptr + self.r0 * ld * val.element_size()], flags=flags | FLAG_STITCH_DIRECT)
            self.stitch = "peer"            # same device-side barrier closes the step
            peer_ptrs = None
        elif self.stitch == "peer":
            esz = val.element_size()
            off = self.r0 * ld * esz
            reps = [p + off for r, p in enumerate(peer_ptrs) if r != self.rank]
            self.opts = make_opts(replicas=reps, flags=flags)
        if self.stitch == "peer":
            # flag words of the device-side barrier that completes the step (sblas_peer_barrier)
            import torch
            self.flags = IpcBuffer(256, val.device)
            self.flag_ptrs = self.flags.open_peers()
            self.flags.tensor(torch.float32, 64).zero_()
            self.status = torch.zeros(1, dtype=torch.int32, device=val.device)
            self.epoch = 0
            torch.cuda.synchronize()
            dist.barrier()          # everybody's flags are zero before the first signal

    def compute(self, xB, out, alpha, beta):
        """Local panel only (the part the roofline is quoted on)."""
        if self.r1 <= self.r0:
            return
        panel = out[self.r0 * self.ld:self.r1 * self.ld]
        if self.k == 1 and self.ld == 1:
            K.spmv(self.rp, self.ci, self.val, xB, panel, alpha, beta, plan=self.plan, opts=self.opts,
                   cols=self.cols)
        else:
            K.spmm(self.rp, self.ci, self.val, self.k, ROW_MAJOR, xB, self.ld, panel, self.ld, alpha,
                   beta, self.cols, plan=self.plan, opts=self.opts)

    def exchange(self, out):
        """Makes every replica complete."""
        dist = _dist()
        if self.stitch == "none":
            return
        if self.stitch == "peer":
            # device-side: signal every peer, wait for every peer; no host synchronisation
            self.epoch += 1
            arr = (C.c_void_p * self.world)(*[C.c_void_p(p) for p in self.flag_ptrs])
            check(lib().sblas_peer_barrier(stream_ptr(), C.c_int(self.rank), C.c_int(self.world),
                                           C.c_void_p(self.flags.ptr.value), arr, C.c_int32(self.epoch),
                                           C.c_void_p(self.status.data_ptr())), "peer_barrier")
            return
        if self.equal:
            dist.all_gather_into_tensor(out[:self.rows_total * self.ld],
                                        out[self.r0 * self.ld:self.r1 * self.ld])
        else:
            works = []
            for r in range(self.world):
                a, b = self.row_begin[r] * self.ld, self.row_begin[r + 1] * self.ld
                if b > a:
                    works.append(dist.broadcast(out[a:b], src=r, async_op=True))
            for w in works:
                w.wait()

    def check(self):
        """Raises if a device-side barrier of this op ever timed out (its status word is sticky): the
        replica of y / C is then INCOMPLETE — a peer was more than SBLAS_BARRIER_TIMEOUT_MS late (first-call
        plan build, a host stall, a profiler serialising the ranks).  Synchronises the device."""
        st = getattr(self, "status", None)
        if st is None:
            return
        bad = int(st.item())
        if bad != 0:
            raise RuntimeError("sblas peer barrier: rank %d timed out waiting for rank %d; the replicated "
                               "output of this and later steps is incomplete" % (self.rank, bad - 1))

    def synchronize(self):
        """Host synchronisation point: waits for the device and verifies the barrier status."""
        import torch
        torch.cuda.synchronize()
        self.check()

    def __call__(self, xB, out, alpha, beta):
        self.compute(xB, out, alpha, beta)
        self.exchange(out)
        # the first step is where a rank can be late (lazy module load, plan tables): verify it before the
        # caller can consume the result; afterwards at synchronize() / check(), and every 64th call
        self._calls = getattr(self, "_calls", 0) + 1
        if self._calls == 1 or self._calls % 64 == 0:
            self.check()
        return out
