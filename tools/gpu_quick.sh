timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/host_time.py 2>&1 | grep host
