#!/usr/bin/env python
"""One small HierMPNEncoder forward + backward per GEMM engine (golden hier_lstm_small and hier_gru_small), meant to be run
under `compute-sanitizer --tool memcheck|racecheck|synccheck` (the backward uses red.global.add.v4.f32 and
atomicMax, the tcgen05 epilogues hand-rolled mbarrier / TMA ordering).  Prints the parity error so a sanitizer run also
proves the instrumented kernels still compute the right thing."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("GGPM_ALLOW_TEST_BACKEND", "0")

import torch                       # noqa: E402
import _checks as C                # noqa: E402
from ggpm_b200 import _lib         # noqa: E402

names = sys.argv[1:] or ["hier_lstm_small", "hier_gru_small"]
for eng in (1, 2, 0):
    _lib.set_gemm_engine(eng)
    for name in names:
        C.check_forward_backward_golden(name, "cuda")
        torch.cuda.synchronize()
        worst = max(v["err"] for v in C.REPORT["golden:%s:engine%d" % (name, eng)].values())
        print("engine %d %s: fwd+bwd ok, worst relative error %.2e, launches so far %d" % (eng, name, worst, _lib.launch_count()))
_lib.set_gemm_engine(0)
