#!/usr/bin/env python
"""Host issue time vs total time per step (cfg2): if the host can run ahead of the device the first is well below the
second; if they are equal the step has a host/stream synchronisation in it or is host bound."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, ggpm_b200
from ggpm_b200 import synth, nnutils
w = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "cfg2"]; H = w["hidden"]
torch.manual_seed(0)
model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, 20, 20, w["dropout"]).cuda().train()
bs = bench.make_batches(w, 2, 1)
db = [nnutils.make_cuda((t, g)) for t, g, _ in bs]
def step(i):
    model.zero_grad(set_to_none=True)
    model(*db[i % 2])[0].sum().backward()
for i in range(4): step(i)
torch.cuda.synchronize()
for rep in range(2):
    t0 = time.perf_counter()
    for i in range(10): step(i)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("host issue time per step %.2f ms, total per step %.2f ms" % ((t1 - t0) * 100, (t2 - t0) * 100))
# pure host cost: how long does issuing take when the device is idle-ish (1 step at a time, sync between)
ts = []
for i in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter(); step(i); ts.append(time.perf_counter() - t0)
print("host time of one step issued into an idle stream: %.2f ms (min of 5)" % (min(ts) * 1e3))
