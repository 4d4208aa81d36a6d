#!/usr/bin/env python
"""GPU timeline of one HierMPNEncoder step (torch.profiler / CUPTI): total kernel time, idle time between kernels inside the
step, per-kernel totals.  Usage: python tools/step_timeline.py [--workload cfg4] > profiles/<name>.txt"""
import argparse, os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
import torch
import bench
import ggpm_b200
from ggpm_b200 import ingest, synth, tensorize

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="cfg4")
a = ap.parse_args()
w = bench.WORKLOADS[a.workload]
H = w["hidden"]
torch.manual_seed(0)
model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), w.get("rnn", "LSTM"), H, H, w["depthT"], w["depthG"], w["dropout"])
for p in model.parameters():
    torch.nn.init.constant_(p, 0) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
model = model.cuda().train()
batches = [ingest.pack(tensorize.tensorize(bench.balanced_batch(7000 + i, w["batch"]))).to_device("cuda") for i in range(2)]

def step(i):
    model.zero_grad(set_to_none=True)
    model(*batches[i % 2])[0].sum().backward()

for i in range(6):
    step(i)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for i in range(2):
        step(i)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
busy, last_end, gaps = 0.0, t0, []
tot = collections.Counter(); cnt = collections.Counter()
for e in ev:
    s, en = e.time_range.start, e.time_range.end
    if s > last_end:
        gaps.append((s - last_end, e.name[:60]))
    busy += max(0.0, en - max(s, last_end))
    last_end = max(last_end, en)
    tot[e.name[:70]] += en - s; cnt[e.name[:70]] += 1
span = t1 - t0
print("# %s: 2 steps, GPU span %.2f ms, busy %.2f ms (%.1f %%), idle %.2f ms in %d gaps" % (a.workload, span / 1e3, busy / 1e3, 100 * busy / span, (span - busy) / 1e3, len(gaps)))
gaps.sort(reverse=True)
print("largest gaps (us, kernel that follows):", [(round(g, 1), n) for g, n in gaps[:12]])
print("gap histogram: >100us %d, 20-100us %d, 5-20us %d, <5us %d" % (sum(g > 100 for g, _ in gaps), sum(20 < g <= 100 for g, _ in gaps), sum(5 < g <= 20 for g, _ in gaps), sum(g <= 5 for g, _ in gaps)))
for k, v in tot.most_common(28):
    print("%-72s %5d %10.1f us %5.1f%%" % (k, cnt[k], v, 100 * v / busy))
