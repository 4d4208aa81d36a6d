# launch list of one timed decode loop of the incremental path (ncu, durations only), after a plain run has passed
set -x
B="python tools/inc_time.py --iters 1 --cpu-batch 0"
$B > gpurun_out/${TAG}_inc_plain.json 2> gpurun_out/${TAG}_inc_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60000 --csv --log-file gpurun_out/${TAG}_inc_launches.csv $B > gpurun_out/${TAG}_inc_ncu.log 2>&1
echo "ncu exit $?"
python - <<PY
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/${TAG}_inc_launches.csv") if l.startswith('"')))
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(",", "")); v = v / 1e3 if r[ui] in ("ns", "nsecond") else v
    name = r[ki].split("(")[0][:70]
    agg[name][0] += 1; agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("kernels %d  total %.1f ms" % (sum(v[0] for v in agg.values()), tot / 1e3))
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print("%6d  %9.1f us  %5.1f%%  avg %6.2f us  %s" % (n, us, 100 * us / tot, us / n, k))
PY
