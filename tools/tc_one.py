#!/usr/bin/env python
"""One tcgen05 GEMM launch of the cfg2 gate-projection shape, for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ggpm_b200 import _lib, functional as Fn
mode = sys.argv[1] if len(sys.argv) > 1 else "plain"
E, Hp = 18965, 256
torch.manual_seed(0)
h = torch.randn(E, Hp, device="cuda"); U4 = torch.randn(4 * Hp, Hp, device="cuda") / 16
out = torch.empty(E, 4 * Hp, device="cuda")
rng = np.random.default_rng(0)
idx = rng.integers(1, E, size=(E, 4)) * (rng.random((E, 4)) < 0.4)
csr = Fn.csr_from_padded(torch.from_numpy(idx).to("cuda"), E)
for _ in range(3):
    if mode == "plain":
        Fn.gemm_nt(h, U4, out=out)
    else:
        Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp)
torch.cuda.synchronize()
print("ok")
