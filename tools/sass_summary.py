#!/usr/bin/env python
"""Per-kernel count of the Blackwell-only SASS mnemonics in libggpm_b200.so (the evidence that the hot GEMMs are
tcgen05 / TMEM / TMA code, /opt/skills/guides/B200_PROFILING.md): UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld /
st), UTCBAR (tcgen05.commit), UTMALDG / UTMASTG / UTMAREDG (bulk tensor load / store / reduce), SYNCS (mbarrier),
REDG (global reductions), plus registers per thread from the ELF.  Usage: python tools/sass_summary.py > profiles/<name>.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ggpm_b200", "libggpm_b200.so")
MNEMONICS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "SYNCS", "REDG", "MUFU"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
counts, arch, cur = collections.OrderedDict(), set(), None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        cur = re.sub(r"\(.*", "", cur.replace("(anonymous namespace)::", "")).replace("void ", "")
        counts.setdefault(cur, collections.Counter())
        continue
    m = re.match(r"\s*arch = (\S+)", line)
    if m:
        arch.add(m.group(1))
    if cur:
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            for k in MNEMONICS:
                if op == k or op.startswith(k + "."):
                    counts[cur][k] += 1
            counts[cur]["_all"] += 1
regs = {}
res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
fn = None
for line in res.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        fn = re.sub(r"\(.*", "", subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
                    .replace("(anonymous namespace)::", "")).replace("void ", "")
    m = re.search(r"REG:(\d+)", line)
    if m and fn:
        regs[fn] = int(m.group(1))
print("# %s  arch %s  (%d kernels)" % (os.path.relpath(LIB, ROOT), ",".join(sorted(arch)), len(counts)))
print("%-74s %5s %7s " % ("kernel", "regs", "instrs") + " ".join("%8s" % k for k in MNEMONICS))
tot = collections.Counter()
for k, c in counts.items():
    tot.update(c)
    print("%-74s %5s %7d " % (k[:74], regs.get(k, "?"), c["_all"]) + " ".join("%8d" % c[m] for m in MNEMONICS))
print("%-74s %5s %7d " % ("TOTAL", "", tot["_all"]) + " ".join("%8d" % tot[m] for m in MNEMONICS))
