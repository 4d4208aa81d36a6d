set -x
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-step"
$B > gpurun_out/s47_plain.log 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s47_launches.csv $B > gpurun_out/s47_ncu1.log 2>&1
python - <<'PY' > gpurun_out/s47_gaps.txt 2>&1
import torch, time, sys
sys.path.insert(0,'.')
import bench, ggpm_b200
from ggpm_b200 import synth, nnutils
from torch.profiler import profile, ProfilerActivity
w = bench.WORKLOADS['cfg2']; H=w['hidden']
torch.manual_seed(0)
model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, 20, 20, 0.1).cuda().train()
bs = bench.make_batches(w, 2, 1)
from ggpm_b200 import ingest
db = [ingest.pack((t,g)).to_device("cuda") for t,g,_ in bs]
def step(i):
    model.zero_grad(set_to_none=True)
    model(*db[i%2])[0].sum().backward()
for i in range(4): step(i)
torch.cuda.synchronize()
t0=time.perf_counter()
for i in range(10): step(i)
t1=time.perf_counter()
torch.cuda.synchronize()
t2=time.perf_counter()
print("host issue time per step %.2f ms, total per step %.2f ms"%((t1-t0)*100,(t2-t0)*100))
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(3): step(i)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
PY
