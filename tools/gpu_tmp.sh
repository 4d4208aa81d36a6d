# launch list of one profiled cfg2 step with the current build (ncu durations only), then its per-kernel summary
set -x
T=${TAG:-tmp}
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-step"
$B > gpurun_out/${T}_plain.log 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/${T}_ncu1.log 2>&1
echo "ncu exit $?"
python - <<PY
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/${T}_launches.csv") if l.startswith('"')))
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(",", "")); v = v / 1e3 if r[ui] in ("ns", "nsecond") else v
    name = r[ki].split("(")[0][:80]
    agg[name][0] += 1; agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("cfg2, one profiled step (forward + backward), current build: kernels %d  total %.2f ms (cold-cache, serialised)" % (sum(v[0] for v in agg.values()), tot / 1e3))
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print("%6d  %9.1f us  %5.1f%%  avg %6.2f us  %s" % (n, us, 100 * us / tot, us / n, k))
PY
