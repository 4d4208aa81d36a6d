set -x
T=${TAG:-r2c}
timeout 1200 python -m pytest tests/test_gpu_parity_full.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${T}_pytest.log
TAG=${T} bash tools/gpu_ncu_r2.sh
