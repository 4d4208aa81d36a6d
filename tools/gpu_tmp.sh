nproc; cat /proc/loadavg; nvidia-smi --query-gpu=index,name,clocks.sm --format=csv,noheader
python tools/host_breakdown.py 2>&1 | grep -E "host ms|C-ABI total"
python tools/host_time.py 2>&1 | grep host
CUDA_VISIBLE_DEVICES=1 python tools/host_time.py 2>&1 | grep host
python bench.py --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench N=1', d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
cat /proc/loadavg
