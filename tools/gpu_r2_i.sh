set -x
T=${TAG:-r2i}
for M in 32768 0 8192; do
for W in cfg2 cfg1 cfg5; do
GGPM_SGATHER_MIN_ROWS=$M timeout 600 python bench.py --workload $W --steps 30 --warmup 5 --quick > gpurun_out/${T}_$M_$W.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_$M_$W.json').read().strip().splitlines()[-1]); print('MINROWS $M $W', round(d['ms_per_step'],3))"
done; done
