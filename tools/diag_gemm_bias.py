"""Is the tensor-core accumulation error of the f16x3 GEMM biased?  (B200)
  * signed mean of (out - ref) against its rms, random-sign data;
  * out(A, W) + out(-A, W): exactly zero for a sign-symmetric (round-toward-zero) truncation, negative for a floor."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                    # noqa: E402
from ggpm_b200 import functional as Fn          # noqa: E402


def gemm(A, W):
    M, K = A.shape
    N = W.shape[0]
    pa, sa = Fn.f16_prepare_weights(A)
    pb, sb = Fn.f16_prepare_weights(W)
    out = torch.empty(M, N, device="cuda")
    Fn.gemm_f16x3_planes(pa, sa, pb, sb, out, M, N, K)
    return out


for M, N, K in [(8192, 512, 512), (8192, 512, 2048)]:
    torch.manual_seed(K)
    A = torch.randn(M, K, device="cuda") * 3e-3
    W = torch.randn(N, K, device="cuda") * 0.07
    ref = A.double() @ W.double().t()
    rms = float(ref.pow(2).mean().sqrt())
    for name, out in (("f16x3", gemm(A, W)), ("torch fp32", A @ W.t())):
        e = out.double() - ref
        print("K %5d %-10s mean(err)/rms(ref) %+.3e   rms(err)/rms(ref) %.3e   max|err|/max|ref| %.3e" % (
            K, name, float(e.mean()) / rms, float(e.pow(2).mean().sqrt()) / rms, float(e.abs().max() / ref.abs().max())))
    s = gemm(A, W).double() + gemm(-A, W).double()
    print("K %5d out(A,W) + out(-A,W): mean %+.3e of rms(ref), fraction of non-zero elements %.3f" % (
        K, float(s.mean()) / rms, float((s != 0).double().mean())))
    Ap, Wp = A.abs(), W.abs()
    refp = Ap.double() @ Wp.double().t()
    for name, out in (("f16x3", gemm(Ap, Wp)), ("torch fp32", Ap @ Wp.t())):
        print("K %5d %-10s all-positive data: mean relative error %+.3e" % (K, name, float(((out.double() - refp) / refp).mean())))
