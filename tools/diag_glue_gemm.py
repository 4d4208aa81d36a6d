"""Generic (glue) GEMM error vs fp64 at the shapes of the hidden-500 glue linears (B200); run under GGPM_GLUE_KSLICE /
GGPM_GLUE_SIMT to compare the engines."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                    # noqa: E402
from ggpm_b200 import functional as Fn          # noqa: E402

tag = " ".join("%s=%s" % (k, os.environ[k]) for k in ("GGPM_GLUE_KSLICE", "GGPM_GLUE_SIMT") if k in os.environ) or "default"
for M, N, K in [(22000, 512, 1000), (22000, 2048, 520), (75000, 512, 536)]:
    torch.manual_seed(K)
    A = torch.relu(torch.randn(M, K, device="cuda"))
    W = torch.randn(N, K, device="cuda") / K ** 0.5
    b = torch.randn(N, device="cuda") * 0.1
    out = Fn.gemm_nt(A, W, bias=b)
    ref = A.double() @ W.double().t() + b.double()
    f32 = A @ W.t() + b
    e, e32 = out.double() - ref, f32.double() - ref
    rms = float(ref.pow(2).mean().sqrt())
    print("%-22s M %6d N %5d K %5d  rms(err)/rms %.2e  max|err|/max|ref| %.2e   | torch fp32 %.2e %.2e" % (
        tag, M, N, K, float(e.pow(2).mean().sqrt()) / rms, float(e.abs().max() / ref.abs().max()),
        float(e32.pow(2).mean().sqrt()) / rms, float(e32.abs().max() / ref.abs().max())))
