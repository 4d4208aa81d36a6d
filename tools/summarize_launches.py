#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name, launches, total and
share of device time.  Usage: python tools/summarize_launches.py launches.csv > profiles/summary.txt"""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    tot, cnt = defaultdict(float), defaultdict(int)
    with open(path, newline="") as f:
        rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
    hdr = rows[0]
    iname, ival, imet = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    for r in rows[1:]:
        if r[imet] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[iname])
        name = re.sub(r"^void ", "", name)
        tot[name] += float(r[ival].replace(",", "")) / 1e3
        cnt[name] += 1
    total = sum(tot.values())
    print("# %s: %d launches, %.1f us of device time (cold-cache, serialised: compare shares)" % (path, sum(cnt.values()), total))
    print("%-90s %8s %12s %8s %7s" % ("kernel", "launches", "total_us", "avg_us", "share"))
    for k in sorted(tot, key=tot.get, reverse=True):
        print("%-90s %8d %12.1f %8.2f %6.1f%%" % (k[:90], cnt[k], tot[k], tot[k] / cnt[k], 100 * tot[k] / total))


if __name__ == "__main__":
    main(sys.argv[1])
