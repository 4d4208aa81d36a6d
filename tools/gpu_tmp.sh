for rep in 1 2 3; do
for per in 0.2 0.05 5; do
GGPM_BENCH_CLOCK_PERIOD=$per timeout 600 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/tmp_b.json 2> gpurun_out/tmp_b.err
echo "period=$per: $(python tools/bench_summary.py gpurun_out/tmp_b.json | grep 'mol') $(python -c "
import json; d=json.loads(open('gpurun_out/tmp_b.json').read().strip().splitlines()[-1]); print(d['clocks']['samples'])")"
done
done
python - <<'PY'
import time, torch, pynvml as nv
nv.nvmlInit(); h=nv.nvmlDeviceGetHandleByIndex(0)
x=torch.randn(8192,8192,device='cuda')
for i in range(20): y=x@x
ts=[]
for i in range(10):
    t0=time.perf_counter(); nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM); t1=time.perf_counter(); nv.nvmlDeviceGetCurrentClocksThrottleReasons(h); t2=time.perf_counter()
    ts.append((round((t1-t0)*1e3,2), round((t2-t1)*1e3,2)))
torch.cuda.synchronize(); print("nvml call ms (clock, reasons):", ts)
PY
