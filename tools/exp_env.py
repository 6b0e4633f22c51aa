#!/usr/bin/env python
"""tools/exp_env.py WORKLOAD VAR=v1,v2,... [VAR2=...] [-- bench args] — runs bench.py (kernel leg only) once per
combination of the given environment variables (the library's experiment hooks) and prints ms/step + frac."""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
if "--" in args:
    i = args.index("--")
    args, extra = args[:i], args[i + 1:]
wl = args[0]
names = [a.split("=")[0] for a in args[1:]]
vals = [a.split("=")[1].split(",") for a in args[1:]]
for combo in itertools.product(*vals):
    env = dict(os.environ)
    for n, v in zip(names, combo):
        if v != "-":
            env[n] = v
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", wl, "--no-e2e", "--no-cpu", "--no-others",
                        "--steps", "10", "--warmup", "3"] + extra, capture_output=True, text=True, env=env, timeout=600)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print(json.dumps(dict(zip(names, combo), ms=d["ms_per_step"], ms_min=d["config"]["ms_per_step_min"], frac=d["roofline"]["frac"],
                              parity=d["parity"]["ok"], err=d["parity"].get("max_err_over_bound"))), flush=True)
    except Exception:
        print("FAILED", combo, r.stdout[-300:], r.stderr[-600:], flush=True)
