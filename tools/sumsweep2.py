#!/usr/bin/env python
"""tools/sumsweep2.py LOG... — one line per sweep point of tools/sweep.py logs."""
import json
import sys
for f in sys.argv[1:]:
    print(f)
    for l in open(f):
        if l.startswith('{'):
            d = json.loads(l)
            print('  ', {k: (round(v, 2) if isinstance(v, float) else v) for k, v in d.items()
                         if k in ('variant', 'flags', 'stages', 'lanes', 'tile', 'ms_med', 'error', 'frac')})
        elif l.startswith('rowblocks'):
            print('  ', l.strip())
