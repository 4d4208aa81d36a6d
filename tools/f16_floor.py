#!/usr/bin/env python
"""Launch a sequence of f16x3 plane GEMMs of known shapes (run under `ncu --metrics gpu__time_duration.sum` to get the
device time of each): fixed cost of a launch, per-chunk time of one CTA, effect of the grid size on a short-K wgrad."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ggpm_b200 import functional as Fn
torch.manual_seed(0)
E, Hp = 5617, 256
Z = torch.randn(E, 4 * Hp, device="cuda"); h = torch.rand(E, Hp, device="cuda"); U4 = torch.randn(4 * Hp, Hp, device="cuda") / 16
pz, sz = Fn.f16_prepare_weights(Z); ph, sh = Fn.f16_prepare_weights(h); pwt, swt = Fn.f16_prepare_weights(U4, transpose=True)
dH = torch.empty(128 * 148, Hp, device="cuda"); dU = torch.zeros(4 * Hp, Hp, device="cuda")
seq = []
def nt(M, N, K, tag):
    Fn.gemm_f16x3_planes(pz, sz, pwt, swt, dH, M, N, K); seq.append("NT M=%d N=%d K=%d %s" % (M, N, K, tag))
def tn(M, N, K, tag, ctas=None):
    if ctas: os.environ["GGPM_F16_CTAS"] = str(ctas)
    Fn.gemm_f16x3_planes(pz, sz, ph, sh, dU, M, N, K, beta=1.0, tn=True); seq.append("TN M=%d N=%d K=%d %s" % (M, N, K, tag))
    os.environ.pop("GGPM_F16_CTAS", None)
for rep in range(2):
    nt(128, 256, 64, "1 tile 1 chunk")
    nt(128, 256, 1024, "1 tile 16 chunks")
    nt(E, 256, 64, "44 tiles 1 chunk")
    nt(E, 256, 1024, "dgrad tree layer")
    tn(128, 256, 64, "1 tile 1 chunk")
    tn(128, 256, 64 * 16, "1 tile 16 chunks, stream-K over 16 CTAs")
    tn(128, 256, 64 * 16, "1 tile 16 chunks, 1 CTA", ctas=1)
    tn(128, 256, 64 * 64, "1 tile 64 chunks, 1 CTA", ctas=1)
    for c in (148, 74, 37, 16, 8):
        tn(1024, 256, E, "wgrad tree layer, %d CTAs" % c, ctas=c)
torch.cuda.synchronize()
print("\n".join(seq))
