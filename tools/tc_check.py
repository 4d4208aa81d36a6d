#!/usr/bin/env python
"""Quick B200 check of the tcgen05 3xTF32 GEMM variants against fp64 (prints errors instead of asserting)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ggpm_b200 import _lib, functional as Fn

DEV = "cuda"
def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30)

torch.manual_seed(0)
_lib.set_gemm_engine(0)
for (M, N, K) in [(128, 128, 32), (128, 256, 32), (128, 128, 256), (129, 257, 252), (1000, 768, 256), (4097, 1024, 512)]:
    A = torch.randn(M, K, device=DEV); W = torch.randn(N, K, device=DEV) / K ** 0.5
    out = Fn.gemm_nt(A, W); torch.cuda.synchronize()
    print("NT  %s rel %.3e" % ((M, N, K), rel(out, A.double() @ W.double().t())), flush=True)
for (M, N, K) in [(128, 128, 32), (300, 252, 768), (1000, 256, 256)]:
    A = torch.randn(M, K, device=DEV); B = torch.randn(K, N, device=DEV) / K ** 0.5
    out = Fn.gemm_nn(A, B); torch.cuda.synchronize()
    print("NN  %s rel %.3e" % ((M, N, K), rel(out, A.double() @ B.double())), flush=True)
for (K, M, N) in [(32, 128, 128), (5000, 768, 256), (16001, 256, 252), (70000, 256, 256)]:
    A = torch.randn(K, M, device=DEV); B = torch.randn(K, N, device=DEV)
    out = Fn.gemm_tn(A, B); torch.cuda.synchronize()
    print("TN  %s rel %.3e" % ((K, M, N), rel(out, A.double().t() @ B.double())), flush=True)
# timing: the LSTM gate GEMM shape of cfg2 / cfg3
for (M, N, K) in [][:0] or [(18965, 1024, 256), (254000, 2048, 512)]:
    A = torch.randn(M, K, device=DEV); W = torch.randn(N, K, device=DEV) / K ** 0.5
    out = torch.empty(M, N, device=DEV)
    for eng in (1, 0):
        _lib.set_gemm_engine(eng)
        Fn.gemm_nt(A, W, out=out); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            Fn.gemm_nt(A, W, out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print("time NT %s engine %d: %.3f ms  %.1f TFLOP/s" % ((M, N, K), eng, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
    _lib.set_gemm_engine(0)
