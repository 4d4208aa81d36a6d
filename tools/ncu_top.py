#!/usr/bin/env python
"""Hottest SASS lines (by warp-stall samples) of one kernel of an .ncu-rep capture.
Usage: python tools/ncu_top.py capture.ncu-rep [ntop] [launch-index]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30; which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
starts.append(len(rows))
blk = rows[starts[which]:starts[which + 1]]
print(blk[0][1][:120])
hdr = blk[1]
data = [r for r in blk[2:] if len(r) == len(hdr) and r[0] != "Address"]
si, so, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
num = lambda x: int(x) if x else 0
print("total samples", sum(num(r[si]) for r in data), "SASS lines", len(data))
agg = {}
for r in data:
    for c in stall:
        agg[hdr[c]] = agg.get(hdr[c], 0) + num(r[c])
print("stall reasons:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
top = sorted(range(len(data)), key=lambda i: -num(data[i][si]))[:ntop]
for i in sorted(top):
    r = data[i]
    st = sorted(((num(r[c]), hdr[c]) for c in stall), reverse=True)[:2]
    print("%5d %6s %8s  %-78s %s" % (i, r[si], r[ie], r[so].strip()[:78], st))
