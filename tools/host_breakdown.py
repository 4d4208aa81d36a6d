#!/usr/bin/env python
"""Where the host time of one cfg2 step goes: seconds spent inside each C-ABI entry point (ctypes call included)
versus everything else (torch ops, autograd, Python)."""
import os, sys, time, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, ggpm_b200
from ggpm_b200 import synth, nnutils, _lib, functional, packing
acc, cnt = collections.Counter(), collections.Counter()
orig = _lib.call
def timed_call(name, *a):
    t0 = time.perf_counter(); orig(name, *a); acc[name] += time.perf_counter() - t0; cnt[name] += 1
_lib.call = functional.call = timed_call
w = bench.WORKLOADS["cfg2"]; H = w["hidden"]
model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, 20, 20, w["dropout"]).cuda().train()
bs = bench.make_batches(w, 2, 1)
db = [nnutils.make_cuda((t, g)) for t, g, _ in bs]
def step(i):
    model.zero_grad(set_to_none=True)
    t0 = time.perf_counter()
    out = model(*db[i % 2])[0].sum()
    t1 = time.perf_counter()
    out.backward()
    return t1 - t0, time.perf_counter() - t1
for i in range(4): step(i)
acc.clear(); cnt.clear()
f = b = 0.0; n = 10
for i in range(n):
    torch.cuda.synchronize(); df, dbk = step(i); f += df; b += dbk
torch.cuda.synchronize()
print("host ms per step: forward %.2f  backward %.2f" % (f / n * 1e3, b / n * 1e3))
tot = 0
for k, v in acc.most_common():
    print("   %-28s %6.3f ms/step  %4d calls/step" % (k, v / n * 1e3, cnt[k] // n)); tot += v
print("   C-ABI total %.2f ms/step; everything else %.2f ms/step" % (tot / n * 1e3, (f + b - tot) / n * 1e3))
