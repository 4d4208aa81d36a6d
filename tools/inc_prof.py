"""host-side profile of one decode loop (cProfile): where the interpreter time of the incremental path goes"""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import inc_time as it                                          # noqa: E402
from ggpm_b200 import encoder as genc                          # noqa: E402
from oracle.ref_loader import FakePairVocab, IdAtomVocab       # noqa: E402

args = type("A", (), dict(kind="hier", rnn="LSTM", hidden=250, diterT=1, diterG=5))()
case = it.make_case(args, 256)
model = genc.IncHierMPNEncoder(FakePairVocab(), IdAtomVocab(), "LSTM", 250, 250, 1, 5, 0.0).to("cuda")
tree_t, graph_t, sched, init = it.inputs(case, "cuda")
init = init.to("cuda").requires_grad_(True)
cots = it.loop(model, True, tree_t, graph_t, sched, init, "cuda")
it.loop(model, True, tree_t, graph_t, sched, init, "cuda", cots)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
it.loop(model, True, tree_t, graph_t, sched, init, "cuda", cots)
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
st.sort_stats("cumulative").print_stats(30)
