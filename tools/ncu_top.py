#!/usr/bin/env python
"""Print headline metrics and the hottest SASS lines (by warp-stall samples) of an .ncu-rep capture."""
import csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
for key in ("gpu__time_duration.sum", "sm__cycles_active.avg", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum"):
    for h, v, u in zip(hdr, vals, rows[1]):
        if h == key:
            print("%-75s %s %s" % (h, v, u))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
si = hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si]) for r in data)
print("total samples", tot)
agg = {}
for r in data:
    for c in stall_cols:
        agg[hdr[c]] = agg.get(hdr[c], 0) + (int(r[c]) if r[c] else 0)
print("stall reasons:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
top = sorted(range(len(data)), key=lambda i: -int(data[i][si]))[:ntop]
for i in sorted(top):
    r = data[i]
    st = sorted(((int(r[c]) if r[c] else 0, hdr[c]) for c in stall_cols), reverse=True)[:2]
    print("%5d %6s  %-70s %s" % (i, r[si], r[1].strip()[:70], st))
