#!/usr/bin/env python
"""Print the headline and the per-kernel-class table of bench.py JSON lines.  Usage: python tools/bench_summary.py f.json ..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e); continue
    print("%s: %.0f mol/s  %.2f ms/step  e2e %.0f  launches %d  alg %.1f TF/s" % (
        f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d.get("algorithmic_tflops", 0)))
    r = d.get("roofline") or {}
    print("   roofline:", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items() if k != "peak_kind"})
    for k, v in (d.get("kernels") or {}).items():
        print("   %-24s %9.3f ms %5d launches %5.1f%%" % (k, v["ms_per_step"], v["launches_per_step"], 100 * v["share"]))
