#!/usr/bin/env python
"""Print the interesting fields of a bench.py JSON line."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.4g %s  ms/step %.4f  n_gpus %s  launches %s" % (d["value"], d["unit"], d["ms_per_step"], d["n_gpus"], d.get("gpu_launches")))
r = d.get("roofline") or {}
print("roofline", {k: r.get(k) for k in ("achieved", "frac", "kernel_ms", "kernel_ms_isolated", "frac_isolated", "traffic")})
print("e2e", d.get("e2e") and {k: d["e2e"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")})
print("cpu", d.get("cpu_baseline") and {k: d["cpu_baseline"][k] for k in ("value", "cores")})
print("clocks", d.get("clocks"), "counts", d.get("counts"))
