#!/usr/bin/env python
"""bench.py — CSR SpMV / SpMM throughput on B200 (GFLOP/s + fraction of the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3v|c4|c5|b32]
                    [--impl ours|reference] [--all]

One "step" = one pass of the hot path (one SpMV / SpMM call) over the synthetic input of a
BASELINE.json config.  At N = 1 the default workload is configs[1] (`c2`: 2-D 5-point
Laplacian, 4M rows, fp64 SpMV).  Prints ONE JSON line (rank 0).  Timing: CUDA events on the
launching stream around every step, L2 flushed between steps, max over ranks.

Workloads (BASELINE.json `configs`):
  c2   Laplacian 2048x2048 grid, fp64 SpMV                         (configs[1])
  c3   uniform 2^24 x 2^24, 32 nnz/row, fp64 SpMM k=64, method 2    (configs[2])
  c3v  the c3 matrix, fp64 SpMV (north-star 16M-row x 32 nnz/row SpMV)
  c4   R-MAT scale 23, edge factor 16, fp32 SpMV                    (configs[3])
  c5   banded 2^21 x 64/row, fp64 SpMM k=256, method 1              (configs[4])
  b32  banded 2^24 x 32/row, fp64 SpMV (structured 16M x 32 counterpart of c3v)
  b32m banded 2^24 x 32/row, fp64 SpMM k=64 (structured 16M x 32 counterpart of c3)
--impl reference times the reference's own CPU implementation (oracle/_ref, else the oracle
port) on the host cores, on a bounded row sample of the same workload.
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

    def __init__(self, index, period=0.004):
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


def run_ours(args, wl_key):
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    K = pkg.kernels
    rank, local_rank, world = env_rank()
    if args.gpus != world and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); "
                         "use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    wl = scaled(WORKLOADS[wl_key], args.scale_down)
    op, f64 = wl["op"], wl["dtype"] == "f64"
    tdtype = torch.float64 if f64 else torch.float32
    esz = 8 if f64 else 4
    k = wl.get("k", 1)
    if args.k and wl["op"] == "spmm":
        k = args.k
    # weak scaling for the SpMV workloads (per-GPU rows fixed), strong for the fixed SpMM configs
    scaling = args.scaling or ("weak" if op == "spmv" else "strong")
    if world > 1 and scaling == "weak":
        if "ny" in wl:
            wl["ny"] *= world
        elif "rows" in wl:
            wl["rows"] *= world
        elif "scale" in wl:
            wl["scale"] += int(np.log2(world))
    rows, cols, rp, ci, v = gen_matrix_gpu(pkg, wl, tdtype, dev)
    nnz = v.numel()
    method = wl.get("method", 2)

    # ---- operands, resident in HBM before the timed region
    sampler = ClockSampler(torch.cuda.current_device() if world == 1 else local_rank)
    sampler.start()
    cfg_extra = {}
    opx = None
    if world == 1:
        plan = K.Plan(rp, nnz)
        if op == "spmm":
            cfg_extra["rowblocks(R, blocks, fill)"] = plan.build_rowblocks(ci, v)   # part of sync2gpu
        ld = k
        xB = pkg.synth.fill_uniform(torch.empty(cols * ld, dtype=tdtype, device=dev), stream_id=11)
        yC = torch.ones(rows * ld, dtype=tdtype, device=dev)
        opts = pkg.make_opts(variant=args.variant, tile=args.tile, lanes=args.lanes, flags=args.flags, stages=args.stages)

        def step_compute():
            if op == "spmv":
                K.spmv(rp, ci, v, xB, yC, 1.0, 1.0, plan=plan, opts=opts, cols=cols)
            else:
                K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, xB, ld, yC, ld, 1.0, 1.0, cols, plan=plan, opts=opts)
        step = step_compute
        loc_rows, loc_nnz, loc_k = rows, nnz, k
        cfg_extra["plan"] = plan.stats()
    elif op == "spmm" and method == 1:
        # method 1 (spmm.h:84-161): A replicated, B/C split by columns, no exchange
        avg, dims = K.dense_split(k, world)
        kk = int(dims[rank])
        plan = K.Plan(rp, nnz)
        cfg_extra["rowblocks(R, blocks, fill)"] = plan.build_rowblocks(ci, v)
        ld = kk
        xB = pkg.synth.fill_uniform(torch.empty(cols * ld, dtype=tdtype, device=dev), stream_id=11 + rank)
        yC = torch.ones(rows * ld, dtype=tdtype, device=dev)

        def step_compute():
            K.spmm(rp, ci, v, kk, pkg.ROW_MAJOR, xB, ld, yC, ld, 1.0, 1.0, cols, plan=plan)
        step = step_compute
        loc_rows, loc_nnz, loc_k = rows, nnz, kk
        cfg_extra["stitch"] = "none (method 1: B/C column split)"
    else:
        # SpMV v1 / SpMM method 2: A split into nnz-balanced row panels, x/B and y/C replicated
        D = pkg.dist
        rb = D.panel_bounds(rp.cpu().numpy(), world)
        r0, r1, lrp, lci, lv = D.local_panel(rp, ci, v, rb, rank)
        lrp = lrp.contiguous()
        lci, lv = lci.clone(), lv.clone()      # 16-byte aligned panel copies
        del rp, ci, v
        torch.cuda.empty_cache()
        ld = k
        xB = pkg.synth.fill_uniform(torch.empty(cols * ld, dtype=tdtype, device=dev), stream_id=11)
        stitch = D.choose_stitch(rb, world) if args.stitch == "auto" else args.stitch
        cfg_extra["stitch_requested"] = args.stitch
        peer_ptrs, ipc = None, None
        mc_ptr = 0
        if stitch == "mc":
            sb = D.SymmBuffer(rows * ld, tdtype, dev)
            yC, peer_ptrs, mc_ptr = sb.tensor(), sb.peer_ptrs, sb.mc_ptr
            yC.fill_(1.0)
            if not mc_ptr:
                stitch = "peer"            # no NVLS on this box: peer stores through the same mapping
            cfg_extra["multicast"] = bool(mc_ptr)
        elif stitch == "peer":
            ipc = D.IpcBuffer(rows * ld * esz, dev)
            peer_ptrs = ipc.open_peers()
            yC = ipc.tensor(tdtype, rows * ld)
            yC.fill_(1.0)
        else:
            yC = torch.ones(rows * ld, dtype=tdtype, device=dev)
        opx = D.RowSplitOp(rows, cols, rb, lrp, lci, lv, k=k, ld=ld if op == "spmm" else 1,
                           stitch=stitch, peer_ptrs=peer_ptrs, mc_ptr=mc_ptr, flags=args.flags)

        def step_compute():
            opx.compute(xB, yC, 1.0, 1.0)

        def step():
            opx.compute(xB, yC, 1.0, 1.0)
            opx.exchange(yC)
        loc_rows, loc_nnz, loc_k = r1 - r0, lv.numel(), k
        cfg_extra["stitch"] = stitch
        # replicated y/C: every GPU must RECEIVE the rows it does not own, whatever the stitch;
        # at NVLink 5's 900 GB/s per direction that alone takes this long per step
        cfg_extra["replication_inbound_bytes_per_gpu"] = int((rows - (r1 - r0)) * (ld if op == "spmm" else 1) * esz)
        cfg_extra["nvlink_floor_ms_per_step"] = cfg_extra["replication_inbound_bytes_per_gpu"] / 900e9 * 1e3
        cfg_extra["panel_rows"] = [int(rb[i + 1] - rb[i]) for i in range(world)]

    flush = torch.empty(2 * L2_BYTES // 4, dtype=torch.int32, device=dev)
    flush_rd = torch.ones_like(flush)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def timed_flushed(fn, steps):
        """per-step device times (CUDA events on the launching stream), L2 flushed between
        steps outside the event pairs"""
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
              for _ in range(steps)]
        for a, b in ev:
            flush.zero_()        # write 252 MiB: evicts every line of the previous step
            flush_rd.sum()       # then read 252 MiB: L2 is left full of CLEAN lines, so the step is
            a.record()           # not billed for the write-back of the flush buffer's dirty lines
            fn()
            b.record()
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in ev]

    def timed_region(fn, steps):
        """exactly `steps` steps back to back between one event pair (inputs larger than L2: every
        step streams its operands from HBM again, no flush needed)"""
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return [a.elapsed_time(b) / steps] * steps

    abytes = algorithmic_bytes(op, loc_rows, cols, loc_nnz, loc_k, esz)   # this rank's launch
    big = abytes > 2 * L2_BYTES        # per-rank operands at least twice the L2: no flush needed
    timed = timed_region if big else timed_flushed

    L = pkg._lib.lib()
    # the very first call (plan tables, kernel attributes, peer mappings are set up lazily): logged
    # because the reference's own timed region includes its per-call setup (spmv.h:43-58)
    ca, cb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ca.record()
    step()
    cb.record()
    torch.cuda.synchronize()
    cfg_extra["cold_call_ms"] = ca.elapsed_time(cb)
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler.active = True
    n0 = L.sblas_launch_count()
    t_wall0 = time.perf_counter()
    ms = timed(step, args.steps)
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    launches = int(L.sblas_launch_count() - n0)
    # kernel-only duration (no exchange) for the roofline; same as `ms` when there is no stitch
    ms_k = ms if step is step_compute else timed(step_compute, args.steps)
    barrier()
    sampler.active = False
    # the other protocol, for information (flushed per-step when the headline is back-to-back and
    # vice versa); single GPU only (per-step events across unsynchronised ranks would time skew)
    ms_other = None
    if world == 1:
        ms_other = float(np.median((timed_flushed if big else timed_region)(step_compute, min(args.steps, 50))))
    if opx is not None and getattr(opx, "status", None) is not None and int(opx.status.item()) != 0:
        raise SystemExit("peer barrier timed out waiting for rank %d" % (int(opx.status.item()) - 1))

    tot_ms = float(np.sum(ms))
    tot_k_ms = float(np.sum(ms_k))
    if dist is not None:
        t = torch.tensor([tot_ms, tot_k_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot_ms, tot_k_ms = float(t[0]), float(t[1])
    flops_step = 2.0 * nnz * k                       # whole job (all ranks)
    value = flops_step * args.steps / (tot_ms * 1e-3) / 1e9
    peak, peak_src = measured_peaks()
    dur_s = tot_k_ms * 1e-3 / args.steps
    achieved = abytes / dur_s / 1e9

    # ---- e2e: the same step through the public call with HOST buffers (pinned): H2D of the
    # step's dense inputs, the op, D2H of the result.  A stays resident (it is distributed once
    # by sync2gpu, outside the reference's own timed region too: spmm_test.cu:28-39).
    e2e = None
    if not args.no_e2e:
        hx = torch.empty(xB.numel(), dtype=tdtype).pin_memory()
        hy = torch.empty(yC.numel(), dtype=tdtype).pin_memory()
        hx.copy_(xB)
        hy.fill_(1.0)
        n_e2e = max(1, min(args.steps, args.e2e_steps))

        def e2e_step():
            xB.copy_(hx, non_blocking=True)
            yC.copy_(hy, non_blocking=True)
            step()
            hy.copy_(yC, non_blocking=True)

        if world == 1:
            # one GPU: the same three phases per step, software-pipelined over three streams and two
            # device buffer sets so that step i's D2H, step i+1's H2D (PCIe is full duplex) and the
            # kernel overlap.  Every step still uploads its own x and y and downloads its own result.
            s_in, s_k, s_out = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
            bufs = [(xB, yC), (torch.empty_like(xB), torch.empty_like(yC))]
            hout = torch.empty(yC.numel(), dtype=tdtype).pin_memory()
            ev_in = [torch.cuda.Event() for _ in bufs]
            ev_done = [torch.cuda.Event() for _ in bufs]
            ev_free = [torch.cuda.Event() for _ in bufs]

            def run_on(xb, yc):
                if op == "spmv":
                    K.spmv(rp, ci, v, xb, yc, 1.0, 1.0, plan=plan, opts=opts, cols=cols)
                else:
                    K.spmm(rp, ci, v, k, pkg.ROW_MAJOR, xb, ld, yc, ld, 1.0, 1.0, cols, plan=plan, opts=opts)

            def e2e_region(n):
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
                        xb.copy_(hx, non_blocking=True)
                        yc.copy_(hy, non_blocking=True)
                        ev_in[i % 2].record()
                    with torch.cuda.stream(s_k):
                        s_k.wait_event(ev_in[i % 2])
                        run_on(xb, yc)
                        ev_done[i % 2].record()
                    with torch.cuda.stream(s_out):
                        s_out.wait_event(ev_done[i % 2])
                        hout.copy_(yc, non_blocking=True)
                        ev_free[i % 2].record()
                for st_ in (s_in, s_k, s_out):
                    main.wait_stream(st_)
                b.record()
                torch.cuda.synchronize()
                return a.elapsed_time(b)
            e2e_region(2)
            barrier()
            ms_e = e2e_region(n_e2e)
            e2e_mode = "pipelined over 3 streams (H2D | kernel | D2H), 2 device buffer sets"
        else:
            e2e_step()
            barrier()
            ms_e = float(np.sum(timed(e2e_step, n_e2e)))
            e2e_mode = "sequential on one stream"
        if dist is not None:
            t = torch.tensor([ms_e], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_e = float(t[0])
        e2e = {"value": flops_step * n_e2e / (ms_e * 1e-3) / 1e9, "unit": "GFLOP/s",
               "h2d_bytes_per_step": int((xB.numel() + yC.numel()) * esz),
               "d2h_bytes_per_step": int(yC.numel() * esz), "steps": n_e2e,
               "ms_per_step": ms_e / n_e2e, "mode": e2e_mode}
    clocks = sampler.stop()

    # ---- CPU baseline beside it (rank 0, N = 1): the reference's CPU function on host cores
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(wl_key, wl, op, f64, k, rp, ci, v, cols, budget_s=args.cpu_seconds)

    traffic, traffic_src, kernel_name = ncu_traffic(wl_key, world, args)
    # second roof of the SpMM workloads (config 5 sits at the FP64 ridge): the box's
    # FP64 multiply-add issue rate, measured here with the library's calibration kernel, in the
    # arithmetic mode the run used (bit-exact = DMUL + DADD per multiply-add, SBLAS_FLAG_FMA = DFMA)
    fp64 = None
    if op == "spmm" and f64:
        fp64 = fp64_roof(L, torch, dev)
        rate = fp64["fma_tmac_per_s"] if (args.flags & pkg.FLAG_FMA) else fp64["mul_add_tmac_per_s"]
        t_fp = loc_nnz * loc_k / (rate * 1e12) * 1e3
        t_hbm = abytes / (peak * 1e9) * 1e3
        fp64.update({"mode": "fma" if (args.flags & pkg.FLAG_FMA) else "mul+add (bit-exact)",
                     "fp64_floor_ms": t_fp, "hbm_floor_ms": t_hbm,
                     "frac_of_binding_roof": max(t_fp, t_hbm) / (tot_k_ms / args.steps)})
    if rank == 0:
        out = {
            "metric": "CSR SpMV/SpMM GFLOP/s and % of HBM roofline at 1/2/4/8 B200 vs reference CPU",
            "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": tot_ms / args.steps,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": wl["dtype"], "data": "synthetic",
            "config": dict({"workload": wl_key + ": " + wl["name"], "rows": rows, "cols": cols,
                            "nnz": nnz, "k": k, "alpha": 1.0, "beta": 1.0, "index": "int32",
                            "timing": ("CUDA events around the K back-to-back steps on the launching stream; no L2 "
                                       "flush: every rank's operands (%.0f MB) are larger than 2x the 126 MB L2" % (abytes / 1e6))
                            if big else
                            ("CUDA events per step on the launching stream; L2 flushed (%d MiB memset + clean "
                             "read sweep of a second buffer) between steps" % (2 * L2_BYTES >> 20)),
                            "ms_per_step_other_protocol": ms_other,
                            "wall_ms_timed_region": wall_ms,
                            "ms_per_step_min": float(np.min(ms)), "ms_per_step_median": float(np.median(ms))},
                           **cfg_extra),
            "clocks": clocks, "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(abytes),
                         "kernel_ms_per_launch": tot_k_ms / args.steps,
                         "frac_of_nominal_8TBs": achieved / 8000.0,
                         "gather_model_bytes_per_launch": int(gather_bytes(op, loc_rows, cols, loc_nnz, loc_k, esz)),
                         "kernel": kernel_name},
        }
        if fp64:
            out["roofline"]["fp64"] = fp64
        if e2e:
            out["e2e"] = e2e
        if cpu:
            out["cpu_baseline"] = cpu
        print(json.dumps(out))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


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


def ncu_traffic(wl_key, world, args):
    """dram__bytes_read.sum + dram__bytes_write.sum of the workload's dominant kernel, per launch,
    from the committed `ncu --set full` capture (profiles/ncu_traffic.json: written from the
    summaries beside it).  Only quoted for the default kernel choice on one GPU; else null."""
    default = not (args.variant or args.tile or args.lanes or args.stages or args.flags) and args.scale_down == 1
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f).get(wl_key)
    except (OSError, ValueError):
        ent = None
    if not ent:
        return None, None, None
    if world != 1 or not default:
        return None, None, ent.get("kernel")
    return ent.get("traffic_bytes"), ent.get("source"), ent.get("kernel")


# ------------------------------------------------------------------------------- CPU arms ----
def _load_cpu_checker():
    """oracle/_ref (the reference's own CPU functions, compiled from its sources) when it was
    built, else the oracle port.  The only place bench.py executes anything under oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    port = pyoracle.Oracle()
    ref = pyoracle.Reference() if pyoracle.Reference.available() else None
    return ref, port


def _time_cpu(fn, budget_s, min_reps=1, max_reps=50):
    fn()                                   # warm caches / page in
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


def _sample_rows(rp_host_full, target_nnz):
    """first R rows holding about target_nnz nnz"""
    R = int(np.searchsorted(rp_host_full, target_nnz, side="left"))
    return max(1, min(R, len(rp_host_full) - 1))


def cpu_baseline(wl_key, wl, op, f64, k, rp, ci, v, cols, budget_s=10.0):
    """The reference's single-threaded CPU function (and the OpenMP port on all cores) timed on
    this box, on a bounded row sample of the SAME matrix (copied down from the device)."""
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    ref, port = _load_cpu_checker()
    chk = ref if ref is not None else port
    dt = np.float64 if f64 else np.float32
    rows = rp.numel() - 1
    rp_h = rp.cpu().numpy()
    nnz = int(rp_h[-1])
    # ~0.4 GFLOP/s single-thread guess, refined by a calibration pass
    target = int(min(nnz, max(2e5, 0.2e9 * budget_s / 2 / max(k, 1))))
    R = _sample_rows(rp_h, target)
    n_s = int(rp_h[R])
    rp_s = np.ascontiguousarray(rp_h[:R + 1])
    ci_s = ci[:n_s].cpu().numpy()
    v_s = v[:n_s].cpu().numpy()
    dev = rp.device
    tdtype = torch.float64 if f64 else torch.float32
    if op == "spmv":
        xB = pkg.synth.fill_uniform(torch.empty(cols, dtype=tdtype, device=dev), stream_id=11).cpu().numpy()
    else:   # col-major host B [cols x k] = transpose of the device row-major operand
        Brm = pkg.synth.fill_uniform(torch.empty(cols * k, dtype=tdtype, device=dev), stream_id=11)
        Bcm = torch.empty(cols * k, dtype=tdtype, device=dev)
        pkg.kernels.transpose(Brm, k, cols, k, Bcm, cols)
        torch.cuda.synchronize()
        xB = Bcm.cpu().numpy()
        del Brm, Bcm
    ts = _time_cpu(_cpu_run(chk, op, rp_s, ci_s, v_s, k, xB, cols), budget_s * 0.6)
    flops = 2.0 * n_s * k
    out = {"value": flops / float(np.median(ts)) / 1e9, "unit": "GFLOP/s", "cores": 1,
           "kind": chk.kind, "sample": "rows [0,%d) of %d (%d of %d nnz), %d passes, median" % (
               R, rows, n_s, nnz, len(ts)),
           "function": "sblas_spmv_csr_cpu spmv.h:16-32" if op == "spmv" else "sblas_spmm_csr_cpu spmm.h:30-70"}
    nthr = port.max_threads()
    ts2 = _time_cpu(_cpu_run(port, op, rp_s, ci_s, v_s, k, xB, cols, threads=nthr), budget_s * 0.4)
    out["omp_port"] = {"value": flops / float(np.median(ts2)) / 1e9, "unit": "GFLOP/s", "cores": nthr,
                       "kind": "port", "passes": len(ts2)}
    return out


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


def run_reference(args, wl_key):
    """--impl reference: the reference's own CPU implementation of the path, host cores only."""
    rank, _, world = env_rank()
    if rank != 0:
        return
    wl = scaled(WORKLOADS[wl_key], args.scale_down)
    op, f64 = wl["op"], wl["dtype"] == "f64"
    dt = np.float64 if f64 else np.float32
    k = wl.get("k", 1)
    ref, port = _load_cpu_checker()
    chk = ref if ref is not None else port
    max_rows = {"spmv": 1 << 22, "spmm": 1 << 14}[op] if k <= 64 else 1 << 12
    rows, cols, rp, ci, v, sample = gen_matrix_cpu(wl, dt, max_rows)
    rng = np.random.default_rng(211)
    xB = rng.random(cols * k).astype(dt)
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
    out = {"impl": "reference",
           "metric": "CSR SpMV/SpMM GFLOP/s and % of HBM roofline at 1/2/4/8 B200 vs reference CPU",
           "value": val, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": len(ts), "warmup": args.warmup,
           "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True,
           "scaling": args.scaling or ("weak" if op == "spmv" else "strong"), "vs_baseline": None,
           "dtype": wl["dtype"], "data": "synthetic",
           "config": {"workload": wl_key + ": " + wl["name"], "k": k, "alpha": 1.0, "beta": 1.0,
                      "sample": sample, "sample_rows": rows, "sample_nnz": n_s},
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
    ap.add_argument("--k", type=int, default=0, help="override the dense width of an SpMM workload (experiments)")
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--scale-down", type=int, default=1, help="shrink the workload (smoke runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=100, help="steps of the end-to-end leg (pipeline fill and drain are inside its timed region)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--ref-seconds", type=float, default=60.0)
    args = ap.parse_args()
    wl_key = args.workload or "c2"
    heavy = WORKLOADS[wl_key]["op"] == "spmm"
    if args.steps is None:
        args.steps = 20 if heavy else 1000
    if args.warmup is None:
        args.warmup = 3 if heavy else 20
    if args.impl == "reference":
        run_reference(args, wl_key)
    else:
        run_ours(args, wl_key)


if __name__ == "__main__":
    main()
