set -x
T=${TAG:-tmp}
for r in 1 2 3; do
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench_$r.json 2> gpurun_out/${T}_bench_$r.err; echo "bench exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_$r.json").read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), {k:(round(v['value']), [round(x,2) for x in v['passes_ms_per_step']]) for k,v in d.items() if k.startswith('e2e')})
PY
done
