#!/usr/bin/env python
"""Phase timestamps (globaltimer, ns) of gemm_f16_kernel CTAs: GGPM_F16_TRACE experiment.
stamps: 0 entry, 1 setup done, 2 first stage full (MMA thread), 3 MMAs issued, 4 accumulators ready (epilogue),
5 epilogue done, 6 after final sync, 7 TMEM freed"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ggpm_b200 import functional as Fn
torch.manual_seed(0)
E, Hp = 5617, 256
Z = torch.randn(E, 4 * Hp, device="cuda"); h = torch.rand(E, Hp, device="cuda"); U4 = torch.randn(4 * Hp, Hp, device="cuda") / 16
pz, sz = Fn.f16_prepare_weights(Z); ph, sh = Fn.f16_prepare_weights(h); pwt, swt = Fn.f16_prepare_weights(U4, transpose=True)
dH = torch.empty(E, Hp, device="cuda"); dU = torch.zeros(4 * Hp, Hp, device="cuda")
dbg = torch.zeros(148, 8, dtype=torch.int64, device="cuda")
def run(name, fn, reps=3):
    for r in range(reps):
        dbg.zero_(); torch.cuda.synchronize()
        os.environ["GGPM_F16_TRACE"] = str(dbg.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        os.environ.pop("GGPM_F16_TRACE")
        torch.cuda.synchronize()
    d = dbg.cpu()
    used = d[:, 0] > 0
    d = d[used]
    t0 = int(d[:, 0].min())
    rel = (d - t0).float() / 1e3
    print("%s: event time %.1f us, %d CTAs; stamps (us from first CTA entry): " % (name, e0.elapsed_time(e1) * 1e3, int(used.sum())))
    names = ["entry", "setup", "1st full", "mma issued", "acc ready", "epi done", "final sync", "tmem free"]
    for i, n in enumerate(names):
        col = rel[:, i][d[:, i] > 0]
        if len(col): print("   %-11s min %7.2f  median %7.2f  max %7.2f" % (n, col.min(), col.median(), col.max()))
run("NT 1 tile 1 chunk", lambda: Fn.gemm_f16x3_planes(pz, sz, pwt, swt, dH, 128, 256, 64))
run("NT dgrad tree layer (E=5617, K=1024)", lambda: Fn.gemm_f16x3_planes(pz, sz, pwt, swt, dH, E, 256, 1024))
run("TN wgrad tree layer (K=5617)", lambda: Fn.gemm_f16x3_planes(pz, sz, ph, sh, dU, 1024, 256, E, beta=1.0, tn=True))
