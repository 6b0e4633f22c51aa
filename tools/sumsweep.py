#!/usr/bin/env python
"""Summarise tools/sweep.py logs: best rows per log."""
import json, sys
for f in sys.argv[1:]:
    rows = [json.loads(l) for l in open(f) if l.startswith("{")]
    ok = sorted([r for r in rows if "ms_med" in r], key=lambda r: r["ms_med"])
    print("==", f, "(%d ok, %d err)" % (len(ok), len(rows) - len(ok)))
    for r in ok[:8]:
        print("  ", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()})
    for r in rows:
        if "error" in r:
            print("  ERR", r)
