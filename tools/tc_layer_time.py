#!/usr/bin/env python
"""Time the three fused tcgen05 launches of one LSTM round (forward gates, state gradient, weight gradient)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ggpm_b200 import _lib, functional as Fn

def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for E, Hp, kbar in [(18965, 256, 1.58), (5617, 256, 1.36), (254000, 512, 1.58)]:
    rng = np.random.default_rng(0)
    h = torch.randn(E, Hp, device="cuda"); U4 = torch.randn(4 * Hp, Hp, device="cuda") / 16
    dA = torch.randn(E, 4 * Hp, device="cuda"); dP = torch.randn(E, Hp, device="cuda")
    out = torch.empty(E, 4 * Hp, device="cuda"); dH = torch.empty(E, Hp, device="cuda"); dU = torch.zeros(4 * Hp, Hp, device="cuda")
    idx = rng.integers(1, E, size=(E, 4)) * (rng.random((E, 4)) < kbar / 4)
    csr = Fn.csr_from_padded(torch.from_numpy(idx).to("cuda"), E); bt = csr.T
    fl = 2.0 * E * 4 * Hp * Hp
    splits = max(1, min(-(-148 // (4 * Hp // 128 * -(-Hp // 256))), -(-E // 256)))
    planes = Fn.tf32_split(U4)
    for tile in (0,):
        tt = timeit(lambda: Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp, tile_n=tile, B_planes=planes))
        tp = timeit(lambda: Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, tile_n=tile, B_planes=planes))
        print('   TMA weights: fwd plain %.3f (%.0f) fwd gather %.3f (%.0f)' % (tp, fl / tp / 1e9, tt, fl / tt / 1e9))
        t1 = timeit(lambda: Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp, tile_n=tile))
        t0 = timeit(lambda: Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, tile_n=tile))
        t2 = timeit(lambda: Fn.gemm_fused(dA, 4 * Hp, 1, U4, Hp, 1, dH, E, Hp, 4 * Hp, a_csr=bt, split_dim=2, split_at=3 * Hp, A2=dP, lda2=Hp, tile_n=tile))
        t3 = timeit(lambda: Fn.gemm_fused(dA, 1, 4 * Hp, h, Hp, 1, dU, 4 * Hp, Hp, E, beta=1.0, b_csr=csr, split_dim=3, split_at=3 * Hp, A2=dP, lda2=Hp, split_k=splits, tile_n=tile))
        print("E=%d Hp=%d tile_n=%d: fwd plain %.3f ms (%.0f TF/s) | fwd gather %.3f (%.0f) | dgrad %.3f (%.0f) | wgrad[%d splits] %.3f (%.0f)"
              % (E, Hp, tile, t0, fl / t0 / 1e9, t1, fl / t1 / 1e9, t2, fl / t2 / 1e9, splits, t3, fl / t3 / 1e9), flush=True)
