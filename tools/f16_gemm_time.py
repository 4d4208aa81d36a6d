#!/usr/bin/env python
"""Time the three f16x3 GEMMs of one LSTM round (gate projection with fused gather, state gradient and weight gradient
from fp16 planes) with CUDA events, warm L2, back to back.  Algorithmic TF/s = 8 E Hp^2 / t."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ggpm_b200 import functional as Fn

def timeit(fn, n=20):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

shapes = [(5617, 256, 1.36), (18965, 256, 1.58), (90401, 512, 1.36), (305609, 512, 1.58)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
for E, Hp, kbar in shapes:
    rng = np.random.default_rng(0)
    h = torch.rand(E, Hp, device="cuda") * 2 - 1
    U4 = torch.randn(4 * Hp, Hp, device="cuda") / 16
    Z = torch.randn(E, 4 * Hp, device="cuda")
    out = torch.empty(E, 4 * Hp, device="cuda"); dH = torch.empty(E, Hp, device="cuda"); dU = torch.zeros(4 * Hp, Hp, device="cuda")
    idx = rng.integers(1, E, size=(E, 4)) * (rng.random((E, 4)) < kbar / 4)
    csr = Fn.csr_from_padded(torch.from_numpy(idx).to("cuda"), E)
    deg = (csr.row_ptr[1:] - csr.row_ptr[:-1]).max().to(torch.int32).reshape(1)
    pw, sw = Fn.f16_prepare_weights(U4)
    pwt, swt = Fn.f16_prepare_weights(U4, transpose=True)
    pz, sz = Fn.f16_prepare_weights(Z)
    ph, sh = Fn.f16_prepare_weights(h)
    fl = 2.0 * E * 4 * Hp * Hp
    t_f = timeit(lambda: Fn.gemm_f16x3(h, Hp, pw, sw, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp, a_deg=deg))
    t_p = timeit(lambda: Fn.gemm_f16x3(h, Hp, pw, sw, out, E, 4 * Hp, Hp))
    t_d = timeit(lambda: Fn.gemm_f16x3_planes(pz, sz, pwt, swt, dH, E, Hp, 4 * Hp))
    t_w = timeit(lambda: Fn.gemm_f16x3_planes(pz, sz, ph, sh, dU, 4 * Hp, Hp, E, beta=1.0, tn=True))
    t_0 = timeit(lambda: Fn.gemm_f16x3_planes(pz, sz, pwt, swt, dH, 128, Hp, 64))
    print("E=%d Hp=%d: fwd+gather %.1f us (%.0f TF/s) | fwd plain %.1f (%.0f) | dgrad planes %.1f (%.0f) | wgrad planes %.1f (%.0f) | "
          "one-tile one-chunk launch %.1f us" % (E, Hp, t_f, fl / t_f / 1e6, t_p, fl / t_p / 1e6, t_d, fl / t_d / 1e6, t_w,
                                               fl / t_w / 1e6, t_0), flush=True)
