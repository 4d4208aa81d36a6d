set -x
T=${TAG:-r2j}
for E in "X=1" "GGPM_F16_BN_FWD=128" "GGPM_F16_BN_WGRAD=128" "GGPM_F16_BN_FWD=128 GGPM_F16_BN_WGRAD=128"; do
for W in cfg2 cfg1 cfg5; do
env $E timeout 600 python bench.py --workload $W --steps 40 --warmup 5 --quick > gpurun_out/${T}_tmp.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_tmp.json').read().strip().splitlines()[-1]); print('KNOB [$E] $W', round(d['ms_per_step'],3))"
done; done
