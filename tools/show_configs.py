#!/usr/bin/env python
"""Prints the `configs` block of a bench.py line (or the output of --configs-only) as a small table."""
import json
import sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
d = d.get("configs", d)
for n, c in d.items():
    print(f"{n:8s} {c['ms_per_launch'] * 1e3:8.2f} us  {c['items_per_s']:.3e} {c['unit']:9s} frac={c['roofline']['frac']:.3f}  {c['launch']}")
