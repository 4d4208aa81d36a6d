set -x
T=${TAG:-r2k}
timeout 1200 python -m pytest tests/test_gpu_encoder.py tests/test_gpu_parity_full.py -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${T}_pytest.log
for NC in 2 1 2 1; do
for W in cfg4 cfg2; do
GGPM_CELL_FWD_NC=$NC timeout 600 python bench.py --workload $W --steps 20 --warmup 5 --pool 4 --quick > gpurun_out/${T}_tmp.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_tmp.json').read().strip().splitlines()[-1]); print('NC $NC $W', round(d['ms_per_step'],3))"
done; done
