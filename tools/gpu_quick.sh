timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "f16x3" 2>&1 | tail -5
python tools/f16_trace.py 2>&1 | grep -v Warn
