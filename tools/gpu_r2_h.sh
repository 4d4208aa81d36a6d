set -x
T=${TAG:-r2h}
for B in 8 4 16 0; do
GGPM_CELL_BPS=$B timeout 600 python bench.py --workload cfg4 --steps 10 --warmup 3 --pool 4 --quick > gpurun_out/${T}_bps$B.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_bps$B.json').read().strip().splitlines()[-1]); print('BPS $B cfg4', round(d['ms_per_step'],2))"
GGPM_CELL_BPS=$B timeout 600 python bench.py --workload cfg2 --steps 30 --warmup 5 --quick > gpurun_out/${T}_bps${B}_cfg2.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_bps${B}_cfg2.json').read().strip().splitlines()[-1]); print('BPS $B cfg2', round(d['ms_per_step'],3))"
done
