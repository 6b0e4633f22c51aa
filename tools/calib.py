#!/usr/bin/env python
"""tools/calib.py — what a plain streaming kernel reaches at the byte counts of the bench
workloads, under the bench's own timing protocol (development aid).

Times torch's copy (read+write) and sum (read only) over `--mb` MB of traffic with
  flush=write       252 MiB memset before each rep (leaves L2 full of DIRTY lines)
  flush=write+read  memset, then a read sweep of a second buffer (L2 full of CLEAN lines)
  flush=none        back to back
"""
import argparse
import json

import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, nargs="+", default=[369.0, 1200.0, 7000.0])
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    F = torch.empty(2 * 126 * 1024 * 1024 // 4, dtype=torch.int32, device=dev)
    G = torch.ones_like(F)
    for mb in args.mb:
        n = int(mb * 1e6 / 8)
        a = torch.ones(n // 2, dtype=torch.float64, device=dev)
        b = torch.empty_like(a)
        c = torch.ones(n, dtype=torch.float64, device=dev)
        for name, fn, traffic in (("copy", lambda: b.copy_(a), 2 * a.numel() * 8), ("sum", lambda: c.sum(), c.numel() * 8)):
            for mode in ("write", "write+read", "none"):
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(args.reps):
                    if mode != "none":
                        F.zero_()
                    if mode == "write+read":
                        G.sum()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    fn()
                    e1.record()
                    torch.cuda.synchronize()
                    ts.append(e0.elapsed_time(e1))
                med = float(np.median(ts))
                print(json.dumps({"kernel": name, "traffic_mb": traffic / 1e6, "flush": mode, "ms_med": med,
                                  "ms_min": float(np.min(ts)), "gbs": traffic / med / 1e6}), flush=True)


if __name__ == "__main__":
    main()
