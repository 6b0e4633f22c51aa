#!/usr/bin/env python
"""bench.py — CSR SpMV / SpMM throughput on B200 (GFLOP/s + fraction of the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c3v|c4|c5|b32|b32m]
                    [--impl ours|reference] [--no-others] [--no-e2e] [--no-cpu]

One "step" = one pass of the hot path (one SpMV / SpMM call) over the synthetic input of a BASELINE.json
config.  The default workload, at every N, is configs[2] (`c3`: uniform-random 2^24 x 2^24, 32 nnz/row, fp64
SpMM k=64, method 2) — the shape the north star's roofline and scaling targets are quoted on and the largest
single-GPU configuration; strong scaling (the matrix is fixed, A is split into row panels over the ranks).
The other BASELINE configs (c2, c3v = the same matrix as SpMV, c4, c5) are measured in the same run with a few
steps each and reported under config.other_workloads.  Prints ONE JSON line (rank 0).

Every timed result is VERIFIED in the same run (the reference's drivers do the same: spmm_test.cu:58-93 runs
the op and the CPU function and prints Validation): `parity` compares rows of the GPU output with the
reference's own CPU function (oracle/_ref, else the oracle port) under the north star's bound
(|got - ref| <= 1e-12 (fp64) / 1e-5 (fp32) * (|alpha| (|A||B|)_ij + |beta C0_ij|)), checks a full-size
linearity property on ALL rows (SpMM: C.1 against an SpMV with B.1; SpMV: a second, independent kernel), and
at N > 1 proves that every replica is bit-identical (checksums all-gathered).  A failed check exits non-zero.

Workloads (BASELINE.json `configs`):
  c3   uniform 2^24 x 2^24, 32 nnz/row, fp64 SpMM k=64, method 2    (configs[2])   <- default
  c2   Laplacian 2048x2048 grid, fp64 SpMV                         (configs[1])
  c3v  the c3 matrix, fp64 SpMV (north-star 16M-row x 32 nnz/row SpMV)
  c4   R-MAT scale 23, edge factor 16, fp32 SpMV                    (configs[3])
  c5   banded 2^21 x 64/row, fp64 SpMM k=256, method 1              (configs[4])
  b32  banded 2^24 x 32/row, fp64 SpMV (structured 16M x 32 counterpart of c3v)
  b32m banded 2^24 x 32/row, fp64 SpMM k=64 (structured 16M x 32 counterpart of c3)
--impl reference times the reference's own CPU implementation (oracle/_ref, else the oracle port) on the host
cores, on a bounded row sample of the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L2_BYTES = 126 * 1024 * 1024
METRIC = "CSR SpMV/SpMM GFLOP/s and % of HBM roofline at 1/2/4/8 B200 vs reference CPU"
NVLINK_MEASURED_GBS = 770.0   # B200_PROFILING.md: measured peer copy per direction (nominal 900)
TOL = {"f64": 1e-12, "f32": 1e-5}


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(
        os.environ.get("WORLD_SIZE", "1"))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------- workloads ----
WORKLOADS = {
    "c2": dict(op="spmv", dtype="f64", gen="laplace2d", nx=2048, ny=2048,
               name="synthetic 2D 5-point Laplacian CSR, 4M rows, fp64 SpMV"),
    "c3": dict(op="spmm", dtype="f64", gen="uniform", rows=1 << 24, deg=32, k=64, method=2,
               name="synthetic uniform-random CSR, 16M rows x 32 nnz/row, fp64 SpMM k=64, method 2"),
    "c3v": dict(op="spmv", dtype="f64", gen="uniform", rows=1 << 24, deg=32,
                name="synthetic uniform-random CSR, 16M rows x 32 nnz/row, fp64 SpMV"),
    "c4": dict(op="spmv", dtype="f32", gen="rmat", scale=23, ef=16,
               name="synthetic R-MAT scale 23 ef 16 CSR, 8M rows, fp32 SpMV"),
    "c5": dict(op="spmm", dtype="f64", gen="banded", rows=1 << 21, bw=64, k=256, method=1,
               name="synthetic banded CSR, 2M rows x 64 nnz/row, fp64 SpMM k=256, method 1"),
    "b32": dict(op="spmv", dtype="f64", gen="banded", rows=1 << 24, bw=32,
                name="synthetic banded CSR, 16M rows x 32 nnz/row, fp64 SpMV"),
    "b32m": dict(op="spmm", dtype="f64", gen="banded", rows=1 << 24, bw=32, k=64, method=2,
                 name="synthetic banded CSR, 16M rows x 32 nnz/row, fp64 SpMM k=64, method 2"),
}
DEFAULT_WORKLOAD = "c3"
# the other BASELINE configs carried in the same line (config.other_workloads); BASELINE.json lists c2 and the
# SpMV of the c3 matrix on one GPU only, c4 and c5 at 1/2/4/8
OTHERS_N1 = ["c2", "c3v", "c4", "c5"]
OTHERS_NX = ["c4", "c5"]


def scaled(wl, scale_down):
    """Smaller instance of a workload (smoke runs / CPU-only sampling)."""
    w = dict(wl)
    if scale_down > 1:
        for key in ("rows",):
            if key in w:
                w[key] = max(w[key] // scale_down, 1024)
        if "nx" in w:
            w["ny"] = max(w["ny"] // scale_down, 8)
        if "scale" in w:
            w["scale"] = max(w["scale"] - int(np.log2(scale_down)), 8)
    return w


def algorithmic_bytes(op, rows, cols, nnz, k, esize):
    """every operand touched once; y/C read and written (alpha=beta=1)."""
    return nnz * (esize + 4) + (rows + 1) * 4 + cols * k * esize + 2 * rows * k * esize


def gather_bytes(op, rows, cols, nnz, k, esize):
    """Secondary model: every nnz fetches its own x element / B row from DRAM."""
    return nnz * (esize + 4) + (rows + 1) * 4 + nnz * k * esize + 2 * rows * k * esize


# --------------------------------------------------------------------------------- clocks ----
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.active = False
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self._halt.is_set():
            try:
                if self.active:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for n, bit in names.items():
                        if r & bit:
                            self.reasons.add(n)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=1.0)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------- GPU arm ----
class Env:
    """torch / package / process-group handles of this rank."""

    def __init__(self, args):
        import torch
        import __graft_entry__ as ge
        self.torch = torch
        self.pkg = ge.load_package()
        self.K = self.pkg.kernels
        self.rank, self.local_rank, self.world = env_rank()
        if args.gpus != self.world and self.world > 1:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, self.world))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); "
                             "use --impl reference for the CPU arm")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.L = self.pkg._lib.lib()
        self.flush = None

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        if self.dist is None:
            return [float(v) for v in vals]
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def flush_l2(self):
        """write 252 MiB, then read a second 252 MiB buffer: L2 is left full of CLEAN foreign lines"""
        torch = self.torch
        if self.flush is None:
            self.flush = (torch.empty(2 * L2_BYTES // 4, dtype=torch.int32, device=self.dev),
                          torch.ones(2 * L2_BYTES // 4, dtype=torch.int32, device=self.dev))
        self.flush[0].zero_()
        self.flush[1].sum()


def gen_matrix_gpu(pkg, wl, tdtype, dev):
    S = pkg.synth
    g = wl["gen"]
    if g == "laplace2d":
        return S.laplace2d(wl["nx"], wl["ny"], tdtype, dev)
    if g == "uniform":
        return S.uniform(wl["rows"], wl["rows"], wl["deg"], tdtype, dev)
    if g == "banded":
        return S.banded(wl["rows"], wl["bw"], tdtype, dev)
    if g == "rmat":
        return S.rmat(wl["scale"], wl["ef"], tdtype, dev)
    raise ValueError(g)


def take_windows(rp, ci, v, rows, n_rows_each, where=(0.0, 0.71)):
    """Host copies of a few row windows of the FULL matrix: [(first_row, rowptr rebased to 0, colidx, val)].
    The second window starts at 71 % of the rows: behind nnz offset 2^30 on the 16M x 32 inputs and, at
    N > 1, inside another rank's panel."""
    out = []
    for f in where:
        a = min(int(rows * f), max(rows - n_rows_each, 0))
        b = min(a + n_rows_each, rows)
        if b <= a or any(a < x[0] + len(x[1]) - 1 and x[0] < b for x in out):
            continue
        rps = rp[a:b + 1].cpu().numpy().astype(np.int64)
        n0, n1 = int(rps[0]), int(rps[-1])
        out.append((a, (rps - n0).astype(np.int32), ci[n0:n1].cpu().numpy(), v[n0:n1].cpu().numpy()))
    return out


class Workload:
    """One BASELINE config resident on this rank: operands, plan, the step closures."""

    def __init__(self, env, args, wl_key, primary):
        torch, pkg, K = env.torch, env.pkg, env.K
        self.env, self.args, self.key, self.primary = env, args, wl_key, primary
        wl = scaled(WORKLOADS[wl_key], args.scale_down)
        self.wl = wl
        self.op, self.f64 = wl["op"], wl["dtype"] == "f64"
        self.tdtype = torch.float64 if self.f64 else torch.float32
        self.esz = 8 if self.f64 else 4
        k = wl.get("k", 1)
        if args.k and self.op == "spmm" and primary:
            k = args.k
        self.k = k
        world, rank, dev = env.world, env.rank, env.dev
        # strong scaling for the BASELINE multi-GPU configs (fixed matrix); --scaling weak grows the SpMV inputs
        self.scaling = (args.scaling if primary and args.scaling else "strong")
        if world > 1 and self.scaling == "weak":
            if "ny" in wl:
                wl["ny"] *= world
            elif "rows" in wl:
                wl["rows"] *= world
            elif "scale" in wl:
                wl["scale"] += int(np.log2(world))
        rows, cols, rp, ci, v = gen_matrix_gpu(pkg, wl, self.tdtype, dev)
        self.rows, self.cols, self.nnz = rows, cols, v.numel()
        self.method = wl.get("method", 2)
        self.cfg = {}
        self.opx = None
        self.ipc = None
        flags = args.flags if primary else 0
        # rows of the full matrix kept on the host for the in-run verification (rank 0 computes the CPU side)
        n_par = args.parity_rows if self.op == "spmv" else max(256, args.parity_rows // max(1, k // 8))
        self.windows = take_windows(rp, ci, v, rows, n_par)
        self.full = (rp, ci, v) if world == 1 else None
        if world == 1:
            self.plan = K.Plan(rp, self.nnz)
            if self.op == "spmm":
                self.cfg["rowblocks(R, blocks, fill)"] = self.plan.build_rowblocks(ci, v)   # part of sync2gpu
            else:
                self.plan.prepare(ci, cols, self.esz)
                self.cfg["colpanels(n, cols)"] = self.plan.build_colpanels(ci, v, cols, n_panels=args.panels if primary else 0)
            self.ld = k
            self.xB = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=self.tdtype, device=dev), stream_id=11)
            self.yC = torch.ones(rows * k, dtype=self.tdtype, device=dev)
            self.opts = pkg.make_opts(variant=args.variant, tile=args.tile, lanes=args.lanes, flags=flags,
                                      stages=args.stages) if primary else None
            self.loc_rows, self.loc_nnz, self.loc_k = rows, self.nnz, k
            self.r0, self.r1 = 0, rows
            self.cfg["plan"] = self.plan.stats()
            self.cfg["stitch"] = "none (one GPU)"
        elif self.op == "spmm" and self.method == 1:
            # method 1 (spmm.h:84-161): A replicated, B/C split by columns, no exchange
            avg, dims = K.dense_split(k, world)
            self.kk = kk = int(dims[rank])
            self.col0 = int(sum(int(d) for d in dims[:rank]))
            self.plan = K.Plan(rp, self.nnz)
            self.cfg["rowblocks(R, blocks, fill)"] = self.plan.build_rowblocks(ci, v)
            self.full = (rp, ci, v)
            self.ld = kk
            # this rank's column slice of the SAME global B (stream 11): B[c, col0 + j] = u(c*k + col0 + j)
            Bfull = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=self.tdtype, device=dev), stream_id=11)
            self.xB = Bfull.view(cols, k)[:, self.col0:self.col0 + kk].contiguous().view(-1)
            del Bfull
            self.yC = torch.ones(rows * kk, dtype=self.tdtype, device=dev)
            self.opts = None
            self.loc_rows, self.loc_nnz, self.loc_k = rows, self.nnz, kk
            self.r0, self.r1 = 0, rows
            self.cfg["stitch"] = "none (method 1: B/C column split)"
        else:
            # SpMV v1 / SpMM method 2: A split into nnz-balanced row panels, x/B and y/C replicated
            D = pkg.dist
            rb = D.panel_bounds(rp.cpu().numpy(), world)
            r0, r1, lrp, lci, lv = D.local_panel(rp, ci, v, rb, rank)
            lrp = lrp.contiguous()
            lci, lv = lci.clone(), lv.clone()      # 16-byte aligned panel copies
            del rp, ci, v
            torch.cuda.empty_cache()
            self.ld = k
            self.xB = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=self.tdtype, device=dev), stream_id=11)
            stitch = D.choose_stitch(rb, world) if args.stitch == "auto" else args.stitch
            self.cfg["stitch_requested"] = args.stitch
            peer_ptrs, mc_ptr = None, 0
            if stitch == "mc":
                sb = D.SymmBuffer(rows * k, self.tdtype, dev)
                self.yC, peer_ptrs, mc_ptr = sb.tensor(), sb.peer_ptrs, sb.mc_ptr
                self.yC.fill_(1.0)
                self.symm = sb
                if not mc_ptr:
                    stitch = "peer"            # no NVLS on this box: peer stores through the same mapping
                self.cfg["multicast"] = bool(mc_ptr)
            elif stitch == "peer":
                self.ipc = D.IpcBuffer(rows * k * self.esz, dev)
                peer_ptrs = self.ipc.open_peers()
                self.yC = self.ipc.tensor(self.tdtype, rows * k)
                self.yC.fill_(1.0)
            else:
                self.yC = torch.ones(rows * k, dtype=self.tdtype, device=dev)
            self.opx = D.RowSplitOp(rows, cols, rb, lrp, lci, lv, k=k, ld=k if self.op == "spmm" else 1,
                                    stitch=stitch, peer_ptrs=peer_ptrs, mc_ptr=mc_ptr, flags=flags)
            self.plan = self.opx.plan
            self.loc_rows, self.loc_nnz, self.loc_k = r1 - r0, lv.numel(), k
            self.r0, self.r1 = r0, r1
            self.cfg["stitch"] = stitch
            # replicated y/C: every GPU must RECEIVE the rows it does not own, whatever the stitch
            inbound = int((rows - (r1 - r0)) * k * self.esz)
            self.cfg["replication_inbound_bytes_per_gpu"] = inbound
            self.cfg["nvlink_floor_ms_per_step"] = {"at_900_GBs_nominal": inbound / 900e9 * 1e3,
                                                    "at_770_GBs_measured_peer_copy": inbound / (NVLINK_MEASURED_GBS * 1e9) * 1e3}
            self.cfg["panel_rows"] = [int(rb[i + 1] - rb[i]) for i in range(world)]
        self.abytes = algorithmic_bytes(self.op, self.loc_rows, cols, self.loc_nnz, self.loc_k, self.esz)
        self.big = self.abytes > 2 * L2_BYTES        # per-rank operands at least twice the L2: no flush needed

    # ---- the step -------------------------------------------------------------------------------
    def compute_on(self, xB, yC):
        K, pkg = self.env.K, self.env.pkg
        if self.opx is not None:
            self.opx.compute(xB, yC, 1.0, 1.0)
        elif self.op == "spmv":
            rp, ci, v = self.full
            K.spmv(rp, ci, v, xB, yC, 1.0, 1.0, plan=self.plan, opts=self.opts, cols=self.cols)
        else:
            rp, ci, v = self.full
            K.spmm(rp, ci, v, self.loc_k, pkg.ROW_MAJOR, xB, self.ld, yC, self.ld, 1.0, 1.0, self.cols,
                   plan=self.plan, opts=self.opts)

    def step_compute(self):
        self.compute_on(self.xB, self.yC)

    def step(self):
        self.compute_on(self.xB, self.yC)
        if self.opx is not None:
            self.opx.exchange(self.yC)

    def has_stitch(self):
        return self.opx is not None

    def close(self):
        torch = self.env.torch
        torch.cuda.synchronize()
        if self.env.dist is not None:
            self.env.dist.barrier()
        if self.opx is not None and getattr(self.opx, "flags", None) is not None:
            self.opx.flags.close()
        self.opx = self.plan = self.full = self.xB = self.yC = None
        if self.ipc is not None:
            self.ipc.close()
            self.ipc = None
        import gc
        gc.collect()
        torch.cuda.empty_cache()


def time_steps(env, w, fn, steps, reps):
    """`reps` timed regions of EXACTLY `steps` steps each, CUDA events on the launching stream, barrier +
    synchronize on both sides of every region.  Inputs at least twice the L2 run back to back (every step
    streams its operands from HBM again); smaller ones get per-step events with the L2 flushed in between,
    outside the event pairs.  Returns per-region milliseconds (max over ranks)."""
    torch = env.torch
    out = []
    for _ in range(reps):
        env.barrier()
        if w.big:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                fn()
            b.record()
            env.barrier()
            ms = a.elapsed_time(b)
        else:
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for a, b in ev:
                env.flush_l2()
                a.record()
                fn()
                b.record()
            env.barrier()
            ms = float(sum(a.elapsed_time(b) for a, b in ev))
        out.append(ms)
    return [env.max_over_ranks([m])[0] for m in out]


# ----------------------------------------------------------------------------- verification ----
def _load_cpu_checker():
    """oracle/_ref (the reference's own CPU functions, compiled from its sources) when it was
    built, else the oracle port.  The only place bench.py executes anything under oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    port = pyoracle.Oracle()
    ref = pyoracle.Reference() if pyoracle.Reference.available() else None
    return ref, port


def host_dense_colmajor(env, w):
    """The dense input as the reference's API holds it on the host: pinned, col-major [cols x k] (k = 1: x).
    Device operands are row-major mirrors (DenseMatrix::sync2gpu converts); this is the inverse move."""
    torch, K = env.torch, env.K
    k = w.loc_k
    h = torch.empty(w.cols * k, dtype=w.tdtype, pin_memory=True)
    if k == 1:
        h.copy_(w.xB)
    else:
        K.dense_download_rows(w.xB, w.ld, w.cols, k, h, w.cols, 0)
    torch.cuda.synchronize()
    return h


def verify(env, w, hB, budget_rows=None):
    """In-run verification of the operands that are timed.  Returns the `parity` object."""
    torch, K, pkg = env.torch, env.K, env.pkg
    ref, port = _load_cpu_checker()
    chk = ref if ref is not None else port
    k, ld, rows, cols = w.loc_k, w.ld, w.rows, w.cols
    tol = TOL[w.wl["dtype"]]
    dt = np.float64 if w.f64 else np.float32
    # one fresh step on C0 = ones through the same call the timed region makes (stitch included)
    w.yC.fill_(1.0)
    env.barrier()
    w.step()
    env.barrier()
    if w.opx is not None:
        w.opx.check()
    par = {"ok": True, "bound": "|got-ref| <= %g * (|alpha| (|A||B|)_ij + |beta C0_ij|)" % tol,
           "against": ("%s CPU function (%s)" % ("reference's own" if chk.kind == "reference" else "oracle port of the reference's",
                                                  "sblas_spmv_csr_cpu spmv.h:16-32" if w.op == "spmv" else "sblas_spmm_csr_cpu spmm.h:30-70"))}
    # (1) row windows against the CPU function (rank 0; every rank holds a full replica except under method 1,
    #     where rank 0 checks its own column slice)
    if (env.rank == 0 or (w.opx is None and env.world > 1)) and hB is not None:
        worst, n_rows, bit = 0.0, 0, True
        B_np = hB.numpy()
        for (a, rps, cis, vs) in w.windows:
            R = len(rps) - 1
            if budget_rows:
                R = min(R, budget_rows)
                rps = rps[:R + 1]
            nn = int(rps[-1])
            cis, vs = cis[:nn], vs[:nn]
            got = w.yC.view(rows, ld)[a:a + R, :k].cpu().numpy()
            if w.op == "spmv":
                if chk.kind == "reference":   # spmv.h:21 asserts square A: pad the window with empty rows
                    rp_sq = np.concatenate([rps, np.full(cols - R, rps[-1], np.int32)])
                    refv = chk.spmv(rp_sq, cis, vs, B_np, np.ones(cols, dt), dt(1), dt(1))[:R]
                else:
                    refv = chk.spmv(rps, cis, vs, B_np, np.ones(R, dt), dt(1), dt(1))
                bound = port.absrow_spmv(rps, cis, vs, B_np) + 1.0
                err = np.abs(got[:, 0].astype(np.float64) - refv.astype(np.float64)) / bound
                bit = bit and np.array_equal(got[:, 0], refv)
            else:
                refC = chk.spmm(rps, cis, vs, k, B_np, np.ones(R * k, dt), dt(1), dt(1)).reshape(k, R)
                err = np.zeros(k)
                for n in range(k):
                    bound = port.absrow_spmv(rps, cis, vs, B_np[n * cols:(n + 1) * cols]) + 1.0
                    err[n] = (np.abs(got[:, n].astype(np.float64) - refC[n].astype(np.float64)) / bound).max()
                bit = bit and np.array_equal(np.ascontiguousarray(got.T), refC)
            worst = max(worst, float(err.max()))
            n_rows += R
        par.update({"checked_rows": n_rows, "windows_first_row": [int(x[0]) for x in w.windows],
                    "max_err_over_bound": worst, "tolerance": tol, "bit_identical_to_cpu": bool(bit)})
        par["ok"] = par["ok"] and worst <= tol
    # (2) a full-size property on ALL rows
    if w.opx is None and env.world == 1:
        rp, ci, v = w.full
        if w.op == "spmm":
            # linearity: (beta*C0 + alpha*A*B).1 = beta*(C0.1) + alpha*A*(B.1): row sums of C against an SpMV
            s_got = w.yC.view(rows, ld)[:, :k].sum(1)
            xs = w.xB.view(cols, ld)[:, :k].sum(1).contiguous()
            ys = torch.full((rows,), float(k), dtype=w.tdtype, device=env.dev)
            K.spmv(rp, ci, v, xs, ys, 1.0, 1.0, cols=cols)
            va = v.abs()
            bnd = torch.full((rows,), float(k), dtype=w.tdtype, device=env.dev)
            K.spmv(rp, ci, va, xs.abs(), bnd, 1.0, 1.0, cols=cols)
            del va
            dev_err = float(((s_got - ys).abs() / bnd).max())
            lim = (2.0 * tol) if w.f64 else (4.0 * tol)
            par["full_size_property"] = {"what": "row sums of C (all %d rows) against beta*(C0.1) + alpha*A*(B.1) by SpMV" % rows,
                                         "max_err_over_bound": dev_err, "tolerance": lim}
            par["ok"] = par["ok"] and dev_err <= lim
        else:
            # a second, independent kernel (sub-warp per row straight from global memory) on all rows
            y2 = torch.ones(rows, dtype=w.tdtype, device=env.dev)
            K.spmv(rp, ci, v, w.xB, y2, 1.0, 1.0, opts=pkg.make_opts(variant=pkg.SPMV_VECTOR), cols=cols)
            va = v.abs()
            bnd = torch.ones(rows, dtype=w.tdtype, device=env.dev)
            K.spmv(rp, ci, va, w.xB.abs(), bnd, 1.0, 1.0, opts=pkg.make_opts(variant=pkg.SPMV_VECTOR), cols=cols)
            del va
            dev_err = float(((w.yC - y2).abs() / bnd).max())
            par["full_size_property"] = {"what": "all %d rows against the independent sub-warp-per-row kernel" % rows,
                                         "max_err_over_bound": dev_err, "tolerance": 2.0 * tol}
            par["ok"] = par["ok"] and dev_err <= 2.0 * tol
    # (3) N > 1, replicated output: every replica bit-identical (64-bit wrap-around sum of the bit patterns)
    if env.dist is not None and w.opx is not None:
        ck = w.yC.view(torch.int64 if w.f64 else torch.int32).to(torch.int64).sum().reshape(1) if not w.f64 else \
            w.yC.view(torch.int64).sum().reshape(1)
        allck = [torch.zeros_like(ck) for _ in range(env.world)]
        env.dist.all_gather(allck, ck)
        same = all(int(c) == int(allck[0]) for c in allck)
        par["replicas_identical"] = bool(same)
        par["replica_checksum"] = "%016x" % (int(allck[0]) & 0xffffffffffffffff)
        par["ok"] = par["ok"] and same
    ok = env.max_over_ranks([0.0 if par["ok"] else 1.0])[0] == 0.0
    par["ok"] = bool(ok)
    return par


# ------------------------------------------------------------------------------------ e2e ----
def e2e_measure(env, w, hB, budget_s, max_steps):
    """The same step through the reference-facing calls with HOST buffers (pinned; dense operands col-major as
    the reference's API holds them): every step uploads its x/B and y/C (DenseMatrix::sync2gpu: PCIe copy +
    layout conversion, sblas_dense_upload_rows_*), runs the op (stitch included) and downloads the result
    (sync2cpu, sblas_dense_download_rows_*).  A stays resident: it is distributed once by sync2gpu, outside the
    reference's own timed region too (spmm_test.cu:28-39).
      N = 1: software-pipelined over three streams and two device buffer sets (upload i+1 | op i | download i-1).
      N > 1, row split: every rank uploads only ITS 1/N row slice of x/B and its own panel of y/C, the slices
             are all-gathered over NVLink (NCCL), the op runs with its stitch, every rank downloads its own panel
             — the host<->device traffic of the whole job is one B + one C in and one C out, spread over N links.
      N > 1, method 1: every rank moves its own column slices; no exchange."""
    torch, K, pkg, dist = env.torch, env.K, env.pkg, env.dist
    k, ld, rows, cols, esz = w.loc_k, w.ld, w.rows, w.cols, w.esz
    world, rank = env.world, env.rank
    spmm = w.op == "spmm"

    def up(h, h_ld, r0, nr, dst, stream=None):      # host col-major rows -> device row-major block
        if spmm:
            K.dense_upload_rows(h, h_ld, r0, nr, k, dst, ld, stream=stream)
        else:
            dst[:nr].copy_(h[r0:r0 + nr], non_blocking=True)

    def down(src, nr, h, h_ld, r0, stream=None):
        if spmm:
            K.dense_download_rows(src, ld, nr, k, h, h_ld, r0, stream=stream)
        else:
            h[r0:r0 + nr].copy_(src[:nr], non_blocking=True)

    if world == 1:
        hC0 = torch.ones(rows * k, dtype=w.tdtype, pin_memory=True)
        hOut = torch.empty(rows * k, dtype=w.tdtype, pin_memory=True)
        s_in, s_k, s_out = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
        bufs = [(w.xB, w.yC), (torch.empty_like(w.xB), torch.empty_like(w.yC))]
        ev_in = [torch.cuda.Event() for _ in bufs]
        ev_done = [torch.cuda.Event() for _ in bufs]
        ev_free = [torch.cuda.Event() for _ in bufs]

        def region(n):
            main = torch.cuda.current_stream()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for st_ in (s_in, s_k, s_out):
                st_.wait_stream(main)
            for i in range(n):
                xb, yc = bufs[i % 2]
                with torch.cuda.stream(s_in):
                    if i >= 2:
                        s_in.wait_event(ev_free[i % 2])
                    up(hB, cols, 0, cols, xb, s_in)
                    up(hC0, rows, 0, rows, yc, s_in)
                    ev_in[i % 2].record()
                with torch.cuda.stream(s_k):
                    s_k.wait_event(ev_in[i % 2])
                    w.compute_on(xb, yc)
                    ev_done[i % 2].record()
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_done[i % 2])
                    down(yc, rows, hOut, rows, 0, s_out)
                    ev_free[i % 2].record()
            for st_ in (s_in, s_k, s_out):
                main.wait_stream(st_)
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b)
        t2 = region(2)
        n = int(max(3, min(max_steps, budget_s * 1e3 / max(t2 / 2, 1e-3))))
        ms = region(n)
        mode = ("pinned col-major host operands; per step: upload x/B + y/C (PCIe + device layout conversion), op, "
                "download; pipelined over 3 streams (H2D | kernel | D2H), 2 device buffer sets")
        h2d, d2h = (cols + rows) * k * esz, rows * k * esz
        # the result of the LAST e2e step, as the host got it, against the device result of the same inputs
        w.yC.fill_(1.0)
        w.step_compute()
        torch.cuda.synchronize()
        chk_rows = min(rows, 4096)
        dv = w.yC.view(rows, ld)[:chk_rows, :k].cpu()
        hv = hOut.view(k, rows)[:, :chk_rows].t() if spmm else hOut[:chk_rows].view(-1, 1)
        host_ok = bool(torch.equal(dv, hv.contiguous()))
        del bufs
    else:
        r0, r1 = (w.r0, w.r1) if w.opx is not None else (0, rows)
        nr = r1 - r0
        main = torch.cuda.current_stream()
        if w.opx is not None:
            # my 1/N row slice of x/B (all-gathered below) and my own panel of y/C
            assert cols % world == 0, "e2e slices: cols must divide by the number of ranks"
            cs = cols // world
            hBs = torch.empty(cs * k, dtype=w.tdtype, pin_memory=True)
            # the slice of the SAME global operand: rows [rank*cs, (rank+1)*cs) of the device mirror
            down(w.xB.view(cols, ld)[rank * cs:(rank + 1) * cs].reshape(-1), cs, hBs, cs, 0)
            hCp = torch.ones(max(nr, 1) * k, dtype=w.tdtype, pin_memory=True)
            hOut = torch.empty(max(nr, 1) * k, dtype=w.tdtype, pin_memory=True)
            torch.cuda.synchronize()
            slice_dst = w.xB.view(cols, ld)[rank * cs:(rank + 1) * cs].reshape(-1)
            panel = w.yC.view(rows, ld)[r0:r1].reshape(-1)

            def one():
                up(hBs, cs, 0, cs, slice_dst)
                dist.all_gather_into_tensor(w.xB, slice_dst)      # in place: my slice is already at its offset
                if nr > 0:
                    up(hCp, nr, 0, nr, panel)
                w.step()
                if nr > 0:
                    down(panel, nr, hOut, nr, 0)
            h2d, d2h = (cs + nr) * k * esz, nr * k * esz
            mode = ("pinned col-major host slices; per step every rank uploads its 1/N row slice of x/B and its own panel "
                    "of y/C, NCCL all-gather of the x/B slices over NVLink, op + stitch, download of its own panel")
        else:
            hBs = torch.empty(cols * k, dtype=w.tdtype, pin_memory=True)
            down(w.xB, cols, hBs, cols, 0)
            hCp = torch.ones(rows * k, dtype=w.tdtype, pin_memory=True)
            hOut = torch.empty(rows * k, dtype=w.tdtype, pin_memory=True)
            torch.cuda.synchronize()

            def one():
                up(hBs, cols, 0, cols, w.xB)
                up(hCp, rows, 0, rows, w.yC)
                w.step()
                down(w.yC, rows, hOut, rows, 0)
            h2d, d2h = (cols + rows) * k * esz, rows * k * esz
            mode = "pinned col-major host column slices (method 1); per step upload B/C slices, op, download; no exchange"

        def region(n):
            env.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                one()
            b.record()
            env.barrier()
            return a.elapsed_time(b)
        t2 = env.max_over_ranks([region(2)])[0]
        n = int(max(3, min(max_steps, budget_s * 1e3 / max(t2 / 2, 1e-3))))
        ms = env.max_over_ranks([region(n)])[0]
        # totals over the job
        tot = env.max_over_ranks([0.0])  # placeholder for symmetry
        h2d_all = torch.tensor([float(h2d), float(d2h)], dtype=torch.float64, device=env.dev)
        dist.all_reduce(h2d_all)
        h2d, d2h = int(h2d_all[0]), int(h2d_all[1])
        host_ok = True
        if w.opx is not None:
            w.opx.check()
    flops_step = 2.0 * w.nnz * w.k
    return {"value": flops_step * n / (ms * 1e-3) / 1e9, "unit": "GFLOP/s",
            "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n,
            "ms_per_step": ms / n, "mode": mode, "host_result_matches_device": host_ok}


# ------------------------------------------------------------------------------- CPU arms ----
def _time_cpu(fn, budget_s, min_reps=1, max_reps=50):
    ts = []
    t_end = time.perf_counter() + budget_s
    while len(ts) < max_reps and (len(ts) < min_reps or time.perf_counter() < t_end):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return ts


def _cpu_run(chk, op, rp, ci, v, k, xB, cols, threads=None):
    """One pass of the CPU implementation over rows [0, len(rp)-1); returns a closure."""
    h = len(rp) - 1
    dt = v.dtype
    if op == "spmv":
        y = np.ones(max(h, cols) if threads is None else h, dt)
        if threads is None:
            if chk.kind == "reference":      # spmv.h:21 asserts square A: pad with empty rows
                rp_sq = np.concatenate([rp, np.full(cols - h, rp[-1], np.int32)]) if cols > h else rp
                return lambda: chk.spmv(rp_sq, ci, v, xB, y, dt.type(1), dt.type(1))
            return lambda: chk.spmv(rp, ci, v, xB, y, dt.type(1), dt.type(1))
        return lambda: chk.spmv_rows_omp(0, h, rp, ci, v, xB, y, dt.type(1), dt.type(1), threads)
    Cm = np.ones(h * k, dt)
    if threads is None:
        return lambda: chk.spmm(rp, ci, v, k, xB, Cm, dt.type(1), dt.type(1))
    return lambda: chk.spmm_rows_omp(0, h, rp, ci, v, k, xB, cols, Cm, h, dt.type(1), dt.type(1), threads)


def cpu_baseline(env, w, hB, budget_s):
    """The reference's single-threaded CPU function (and the OpenMP port on all cores) timed on this box, on a
    bounded row sample of the SAME matrix and the same dense operand (copied down from the device)."""
    ref, port = _load_cpu_checker()
    chk = ref if ref is not None else port
    rp, ci, v = w.full
    k, cols, rows, nnz = w.loc_k, w.cols, w.rows, w.nnz
    B_np = hB.numpy()
    # calibrate on a small prefix, then size the sample for ~60 % of the budget
    R0 = min(rows, 2048 if w.op == "spmm" else 1 << 18)

    def sample(R):
        rps = rp[:R + 1].cpu().numpy()
        n = int(rps[-1])
        return rps, ci[:n].cpu().numpy(), v[:n].cpu().numpy(), n
    rps, cis, vs, n0 = sample(R0)
    t0 = min(_time_cpu(_cpu_run(chk, w.op, rps, cis, vs, k, B_np, cols), 0.0, min_reps=2))
    R = int(min(rows, max(R0, R0 * 0.6 * budget_s / max(t0, 1e-6))))
    rps, cis, vs, n_s = sample(R)
    ts = _time_cpu(_cpu_run(chk, w.op, rps, cis, vs, k, B_np, cols), 0.0, min_reps=1)
    flops = 2.0 * n_s * k
    out = {"value": flops / float(np.median(ts)) / 1e9, "unit": "GFLOP/s", "cores": 1,
           "kind": chk.kind, "sample": "rows [0,%d) of %d (%d of %d nnz), %d pass(es) of %.1f s" % (
               R, rows, n_s, nnz, len(ts), float(np.median(ts))),
           "function": "sblas_spmv_csr_cpu spmv.h:16-32" if w.op == "spmv" else "sblas_spmm_csr_cpu spmm.h:30-70"}
    nthr = port.max_threads()
    ts2 = _time_cpu(_cpu_run(port, w.op, rps, cis, vs, k, B_np, cols, threads=nthr), 0.25 * budget_s)
    out["omp_port"] = {"value": flops / float(np.median(ts2)) / 1e9, "unit": "GFLOP/s", "cores": nthr,
                       "kind": "port", "passes": len(ts2)}
    return out


def fp64_roof(L, torch, dev):
    """FP64 multiply-adds per second of this GPU: DFMA chains and DMUL+DADD chains (sblas_fp64_rate)."""
    import ctypes as C
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    grid, iters = 148 * 8, 4096
    macs = grid * 256 * 8 * iters
    res = {}
    for mode, key in ((0, "fma_tmac_per_s"), (1, "mul_add_tmac_per_s")):
        best = None
        for _ in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            rc = L.sblas_fp64_rate(C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_int32(mode),
                                   C.c_int32(grid), C.c_int32(iters), C.c_void_p(out.data_ptr()))
            b.record()
            torch.cuda.synchronize()
            if rc != 0:
                raise SystemExit("sblas_fp64_rate failed")
            t = a.elapsed_time(b)
            best = t if best is None else min(best, t)
        res[key] = macs / (best * 1e-3) / 1e12
    return res


def ncu_traffic(wl_key, world, default_choice):
    """dram__bytes_read.sum + dram__bytes_write.sum of the workload's kernels PER STEP (all launches of a step
    summed), from the committed `ncu --set full` capture (profiles/ncu_traffic.json: written from the summaries
    beside it).  Only quoted for the default kernel choice on one GPU; else null."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f).get(wl_key)
    except (OSError, ValueError):
        ent = None
    if not ent:
        return None, None, None
    if world != 1 or not default_choice or ent.get("traffic_bytes") is None:
        return None, None, ent.get("kernel")
    return int(ent["traffic_bytes"]) * int(ent.get("launches_per_step", 1)), ent.get("source"), ent.get("kernel")


def measure(env, args, wl_key, primary, sampler=None):
    """Build, time and verify one workload; returns the result dict (rank 0's view)."""
    torch, pkg, L = env.torch, env.pkg, env.L
    w = Workload(env, args, wl_key, primary)
    steps = args.steps if primary else (5 if w.op == "spmm" else 200)
    warm = max(args.warmup, 3) if primary else 3
    res = {}
    # the very first call (lazily created tables, kernel attributes, peer mappings): logged because the
    # reference's own timed region includes its per-call setup (spmv.h:43-58)
    ca, cb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ca.record()
    w.step()
    cb.record()
    torch.cuda.synchronize()
    w.cfg["cold_call_ms"] = ca.elapsed_time(cb)
    if w.opx is not None:
        w.opx.check()
    for _ in range(warm):
        w.step()
    env.barrier()
    # number of timed regions of K steps: >= 5 when a region is shorter than 50 ms, 3 below 2 s, else 1 (min / median reported)
    t_probe = time_steps(env, w, w.step, max(1, min(steps, 3)), 1)[0] / max(1, min(steps, 3))
    region_ms = t_probe * steps
    reps = int(min(25, max(5, np.ceil(250.0 / max(region_ms, 1e-3))))) if region_ms < 50.0 else (3 if region_ms < 2000.0 else 1)
    if sampler is not None:
        sampler.active = True
    n0 = L.sblas_launch_count()
    t_wall0 = time.perf_counter()
    regions = time_steps(env, w, w.step, steps, reps)
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    launches = int(L.sblas_launch_count() - n0) // reps      # launches of one timed region (K steps)
    # kernel-only duration (no exchange) for the roofline; same as the step when there is no stitch
    regions_k = regions if not w.has_stitch() else time_steps(env, w, w.step_compute, steps, reps)
    if sampler is not None:
        sampler.active = False
    if w.opx is not None:
        w.opx.check()
    ms_step = float(np.median(regions)) / steps
    ms_k = float(np.median(regions_k)) / steps
    flops_step = 2.0 * w.nnz * w.k                       # whole job (all ranks)
    peak, peak_src = measured_peaks()
    achieved = w.abytes / (ms_k * 1e-3) / 1e9
    gbytes = gather_bytes(w.op, w.loc_rows, w.cols, w.loc_nnz, w.loc_k, w.esz)
    default_choice = not (primary and (args.variant or args.tile or args.lanes or args.stages or args.flags or args.panels)) \
        and args.scale_down == 1
    traffic, traffic_src, kernel_name = ncu_traffic(wl_key, env.world, default_choice)
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "algorithmic_bytes_per_step": int(w.abytes), "kernel_ms_per_step": ms_k,
            "launches_per_step": launches // max(steps, 1), "frac_of_nominal_8TBs": achieved / 8000.0,
            "gather_model_bytes_per_step": int(gbytes),
            "frac_gather_model": gbytes / (ms_k * 1e-3) / 1e9 / peak,
            "kernel": kernel_name}
    if w.op == "spmm" and w.f64 and primary:
        # second roof of the SpMM workloads (config 5 sits at the FP64 ridge): the box's FP64
        # multiply-add issue rate in the arithmetic mode of the run (default: DFMA; SBLAS_FLAG_EXACT: DMUL + DADD)
        fp64 = fp64_roof(L, torch, env.dev)
        exact = bool(args.flags & pkg.FLAG_EXACT)
        rate = fp64["mul_add_tmac_per_s"] if exact else fp64["fma_tmac_per_s"]
        t_fp = w.loc_nnz * w.loc_k / (rate * 1e12) * 1e3
        t_hbm = w.abytes / (peak * 1e9) * 1e3
        fp64.update({"mode": "mul+add (bit-exact)" if exact else "fma (default)", "fp64_floor_ms": t_fp,
                     "hbm_floor_ms": t_hbm, "frac_of_binding_roof": max(t_fp, t_hbm) / ms_k})
        roof["fp64"] = fp64
    res.update({"workload": wl_key + ": " + w.wl["name"], "value": flops_step / (ms_step * 1e-3) / 1e9,
                "ms_per_step": ms_step, "steps": steps, "timed_regions": reps,
                "ms_per_step_min": float(np.min(regions)) / steps, "ms_per_step_median": ms_step,
                "wall_ms_timed": wall_ms, "gpu_launches": launches, "roofline": roof,
                "scaling": w.scaling, "dtype": w.wl["dtype"]})
    if w.has_stitch():
        inbound = w.cfg["replication_inbound_bytes_per_gpu"]
        res["stitch_inbound_gbs"] = inbound / (ms_step * 1e-3) / 1e9
        res["compute_only_ms_per_step"] = ms_k
    # ---- verification of what was timed, e2e, CPU baseline ------------------------------------------------
    need_host = (not args.no_parity) or (primary and not args.no_e2e) or (primary and not args.no_cpu)
    hB = None
    if need_host and (env.rank == 0 or (w.opx is None and env.world > 1)):
        hB = host_dense_colmajor(env, w)
    if not args.no_parity:
        res["parity"] = verify(env, w, hB)
    if primary and not args.no_cpu and env.rank == 0 and env.world == 1:
        res["cpu_baseline"] = cpu_baseline(env, w, hB, args.cpu_seconds)
    if primary and not args.no_e2e:
        res["e2e"] = e2e_measure(env, w, hB if env.world == 1 else None, args.e2e_seconds, args.e2e_steps)
    res["config_extra"] = dict(w.cfg, rows=w.rows, cols=w.cols, nnz=w.nnz, k=w.k,
                               timing=("CUDA events around %d region(s) of K back-to-back steps on the launching stream; no L2 flush: "
                                       "every rank's operands (%.0f MB) are larger than 2x the 126 MB L2" % (reps, w.abytes / 1e6))
                               if w.big else
                               ("CUDA events per step on the launching stream, %d region(s) of K steps; L2 flushed (%d MiB memset + clean "
                                "read sweep of a second buffer) between steps, outside the event pairs" % (reps, 2 * L2_BYTES >> 20)))
    del hB
    w.close()
    return res


def run_ours(args):
    env = Env(args)
    wl_key = args.workload or DEFAULT_WORKLOAD
    sampler = ClockSampler(env.torch.cuda.current_device() if env.world == 1 else env.local_rank)
    sampler.start()
    main = measure(env, args, wl_key, True, sampler)
    clocks = sampler.stop()
    others = []
    if not args.no_others and args.scale_down == 1:
        for key in (OTHERS_N1 if env.world == 1 else OTHERS_NX):
            if key == wl_key:
                continue
            r = measure(env, args, key, False)
            others.append({"workload": r["workload"], "value": r["value"], "unit": "GFLOP/s", "ms_per_step": r["ms_per_step"],
                           "ms_per_step_min": r["ms_per_step_min"], "steps": r["steps"], "timed_regions": r["timed_regions"],
                           "scaling": r["scaling"], "dtype": r["dtype"],
                           "frac": r["roofline"]["frac"], "frac_gather_model": r["roofline"]["frac_gather_model"],
                           "achieved_gbs": r["roofline"]["achieved"], "traffic": r["roofline"]["traffic"],
                           "kernel_ms_per_step": r["roofline"]["kernel_ms_per_step"],
                           "stitch": r["config_extra"].get("stitch"),
                           "colpanels(n, cols)": r["config_extra"].get("colpanels(n, cols)"),
                           "rowblocks(R, blocks, fill)": r["config_extra"].get("rowblocks(R, blocks, fill)"),
                           "parity": r.get("parity")})
    ok = main.get("parity", {}).get("ok", True) and all((o.get("parity") or {}).get("ok", True) for o in others)
    if env.rank == 0:
        cfg = dict({"workload": main["workload"], "alpha": 1.0, "beta": 1.0, "index": "int32",
                    "arithmetic": "fp64 fused multiply-add (SpMM default; SBLAS_FLAG_EXACT = separately rounded)" if main["dtype"] == "f64" else "fp32",
                    "ms_per_step_min": main["ms_per_step_min"], "ms_per_step_median": main["ms_per_step_median"],
                    "timed_regions": main["timed_regions"], "wall_ms_timed_region": main["wall_ms_timed"]},
                   **main["config_extra"])
        for key in ("stitch_inbound_gbs", "compute_only_ms_per_step"):
            if key in main:
                cfg[key] = main[key]
        if others:
            cfg["other_workloads"] = others
        out = {"metric": METRIC, "value": main["value"], "unit": "GFLOP/s", "n_gpus": env.world, "steps": args.steps,
               "warmup": max(args.warmup, 3), "ms_per_step": main["ms_per_step"], "higher_is_better": True,
               "scaling": main["scaling"], "vs_baseline": None, "dtype": main["dtype"], "data": "synthetic",
               "config": cfg, "clocks": clocks, "gpu_launches": main["gpu_launches"], "roofline": main["roofline"]}
        for key in ("parity", "e2e", "cpu_baseline"):
            if key in main:
                out[key] = main[key]
        print(json.dumps(out))
        sys.stdout.flush()
    if env.dist is not None:
        env.dist.barrier()
        env.dist.destroy_process_group()
    if not ok:
        raise SystemExit("bench.py: in-run verification FAILED (see the parity objects of the JSON line)")


def gen_matrix_cpu(wl, dtype, max_rows):
    """numpy inputs for the reference arm (no CUDA): a bounded row sample of the workload."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sblas_synth_np", os.path.join(ROOT, "s-blas_b200", "synth_np.py"))
    SN = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(SN)
    g = wl["gen"]
    if g == "laplace2d":
        rp, ci, v = SN.laplace2d(wl["nx"], wl["ny"], dtype)
        n = wl["nx"] * wl["ny"]
        return n, n, rp, ci, v, "full matrix"
    if g == "banded":
        n, bw = wl["rows"], wl["bw"]
        R = min(n, max_rows)
        c0 = np.clip(np.arange(R) - bw // 2, 0, n - bw)
        ci = (c0[:, None] + np.arange(bw)[None, :]).astype(np.int32).reshape(-1)
        rp = (np.arange(R + 1, dtype=np.int64) * bw).astype(np.int32)
        v = SN.fill_uniform(R * bw, 211, 1, -1.0, 1.0, dtype)
        return R, n, rp, ci, v, "rows [0,%d) of %d" % (R, n)
    if g == "uniform":
        n, deg = wl["rows"], wl["deg"]
        R = min(n, max_rows)
        # same distribution (deg distinct uniform columns per row, sorted); vectorised draw
        rng = np.random.default_rng(211)
        ci = np.sort(rng.integers(0, n, (R, deg), dtype=np.int64), axis=1).astype(np.int32).reshape(-1)
        rp = (np.arange(R + 1, dtype=np.int64) * deg).astype(np.int32)
        v = SN.fill_uniform(R * deg, 211, 1, -1.0, 1.0, dtype)
        return R, n, rp, ci, v, "rows [0,%d) of %d (columns drawn with numpy PCG64, same distribution)" % (R, n)
    if g == "rmat":
        sc = min(wl["scale"], 18)
        rp, ci, v = SN.rmat(sc, wl["ef"], 211, dtype)
        return 1 << sc, 1 << sc, rp, ci, v, "R-MAT scale %d (config: scale %d), same a,b,c,d and edge factor" % (sc, wl["scale"])
    raise ValueError(g)


def _fill_random_parallel(n, dt, threads=8):
    """uniform[0,1) host array of n elements, filled by a few threads (numpy's generators release the GIL)"""
    from concurrent.futures import ThreadPoolExecutor
    out = np.empty(n, dt)
    cuts = np.linspace(0, n, threads + 1).astype(np.int64)

    def fill(i):
        rng = np.random.default_rng(211 + i)
        a, b = int(cuts[i]), int(cuts[i + 1])
        step = 1 << 24
        for s in range(a, b, step):
            e = min(b, s + step)
            if dt == np.float64:
                rng.random(e - s, out=out[s:e])
            else:
                out[s:e] = rng.random(e - s, dtype=np.float32)
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(fill, range(threads)))
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path, host cores only, on a bounded row
    sample of the arm's workload (same matrix family, same dense operand size: the strided gathers into the
    full-size col-major B are what the CPU function spends its time on)."""
    rank, _, world = env_rank()
    if rank != 0:
        return
    wl_key = args.workload or DEFAULT_WORKLOAD
    wl = scaled(WORKLOADS[wl_key], args.scale_down)
    op, f64 = wl["op"], wl["dtype"] == "f64"
    dt = np.float64 if f64 else np.float32
    k = wl.get("k", 1)
    ref, port = _load_cpu_checker()
    chk = ref if ref is not None else port
    max_rows = {"spmv": 1 << 22, "spmm": 1 << 15}[op] if k <= 64 else 1 << 13
    rows, cols, rp, ci, v, sample = gen_matrix_cpu(wl, dt, max_rows)
    xB = _fill_random_parallel(cols * k, dt)
    # size the sample so that one step lasts about ref_seconds / (steps + warmup), at most max_rows rows
    R0 = min(rows, 1024 if op == "spmm" else 1 << 16)
    n0 = int(rp[R0])
    t0 = min(_time_cpu(_cpu_run(chk, op, rp[:R0 + 1].copy(), ci[:n0], v[:n0], k, xB, cols), 0.0, min_reps=2))
    per_step = args.ref_seconds / max(1, args.steps + min(args.warmup, 2))
    R = int(min(rows, max(R0, R0 * per_step / max(t0, 1e-6))))
    if R < rows and wl["gen"] != "laplace2d":
        n_s = int(rp[R])
        rp, ci, v, rows = rp[:R + 1].copy(), ci[:n_s], v[:n_s], R
        sample = sample.replace("rows [0,%d)" % max_rows, "rows [0,%d)" % R) if "rows [0," in sample else sample
    fn = _cpu_run(chk, op, rp, ci, v, k, xB, cols)
    for _ in range(max(1, min(args.warmup, 2))):
        fn()
    ts = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
        if sum(ts) > args.ref_seconds:
            break
    n_s = int(rp[-1])
    val = 2.0 * n_s * k * len(ts) / sum(ts) / 1e9
    out = {"impl": "reference", "metric": METRIC,
           "value": val, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": len(ts), "warmup": args.warmup,
           "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True,
           "scaling": args.scaling or "strong", "vs_baseline": None,
           "dtype": wl["dtype"], "data": "synthetic",
           "config": {"workload": wl_key + ": " + wl["name"], "k": k, "alpha": 1.0, "beta": 1.0,
                      "sample": sample, "sample_rows": int(len(rp) - 1), "sample_nnz": n_s, "dense_operand_elements": int(cols * k)},
           "cpu_baseline": {"value": val, "unit": "GFLOP/s", "cores": 1, "kind": chk.kind, "sample": sample,
                            "function": "sblas_spmv_csr_cpu spmv.h:16-32 (single-threaded in the reference)"
                            if op == "spmv" else "sblas_spmm_csr_cpu spmm.h:30-70 (single-threaded in the reference)"},
           "e2e": {"value": val, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"])
    ap.add_argument("--stitch", default="auto", choices=["auto", "peer", "nccl", "mc"],
                    help="auto: peer stores, multicast stores when the row panels are very unequal (dist.choose_stitch)")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--tile", type=int, default=0)
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--stages", type=int, default=0)
    ap.add_argument("--panels", type=int, default=0, help="column panels of the SpMV mirror (0 = automatic)")
    ap.add_argument("--k", type=int, default=0, help="override the dense width of an SpMM workload (experiments)")
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--scale-down", type=int, default=1, help="shrink the workload (smoke runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip config.other_workloads")
    ap.add_argument("--parity-rows", type=int, default=16384, help="rows per verification window (SpMV; SpMM: / (k/8))")
    ap.add_argument("--e2e-steps", type=int, default=100, help="most steps of the end-to-end leg")
    ap.add_argument("--e2e-seconds", type=float, default=3.0, help="time budget that sizes the end-to-end leg")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--ref-seconds", type=float, default=60.0)
    args = ap.parse_args()
    wl_key = args.workload or DEFAULT_WORKLOAD
    heavy = WORKLOADS[wl_key]["op"] == "spmm"
    if args.steps is None:
        args.steps = 20 if heavy else 1000
    if args.warmup is None:
        args.warmup = 3 if heavy else 20
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
