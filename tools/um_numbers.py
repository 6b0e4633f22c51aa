#!/usr/bin/env python
"""tools/um_numbers.py — SURVEY 8(f) f2 with numbers: the managed-memory drivers (drivers/bench_um, built on
include/matrix_um.h) on c2-size and 16M-row inputs, tuning 0 (pages fault inside the kernels) against tuning 1
(matrix_um.h:529-614: cudaMemAdvise + cudaMemPrefetchAsync before the op).  Prints one JSON line per run and a table."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "drivers", "bin", "bench_um")
RUNS = [("spmv", "c2", 1), ("spmv", "b32", 1), ("spmv", "c3v", 1), ("spmm2", "b32m", 1)]
if len(sys.argv) > 1:
    RUNS = [tuple(a.split(":")[:2]) + (int(a.split(":")[2]),) for a in sys.argv[1:]]
rows = []
for op, wl, n in RUNS:
    for tuning in (0, 1):
        r = subprocess.run([EXE, op, wl, str(n), str(tuning), "3"], capture_output=True, text=True, timeout=900)
        line = [l for l in r.stdout.splitlines() if l.startswith("UM_JSON ")]
        if r.returncode != 0 or not line:
            print("FAILED", op, wl, n, tuning, r.stdout[-500:], r.stderr[-500:], flush=True)
            continue
        d = json.loads(line[0][8:])
        rows.append(d)
        print(json.dumps(d), flush=True)
print("\n| op | workload | GPUs | tuning | hooks before (ms) | first call (ms) | later calls (ms) | host check after 1st call (ms) | hooks after (ms) |")
print("|---|---|---|---|---|---|---|---|---|")
for d in rows:
    print("| %s | %s | %d | %d | %.1f | %.2f | %.3f | %.1f | %.1f |" % (
        d["op"], d["workload"], d["n_gpu"], d["tuning"], d["hooks_before_ms"], d["first_call_ms"], d["later_calls_median_ms"],
        d["host_check_after_first_call_ms"], d["hooks_after_ms"]))
