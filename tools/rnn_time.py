"""Step time of the Hier encoder with either cell at the cfg2 shape (CUDA events, 10 steps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggpm_b200
from ggpm_b200 import ingest, synth, tensorize
B, H = int(os.environ.get("B", 256)), int(os.environ.get("H", 250))
batches = [ingest.pack(tensorize.tensorize(synth.random_batch(5 + i, B))).to_device("cuda") for i in range(3)]
for rnn in ("LSTM", "GRU"):
    torch.manual_seed(0)
    m = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), rnn, H, H, 20, 20, 0.0).cuda()
    for p in m.parameters():
        torch.nn.init.normal_(p, std=0.02) if p.dim() > 1 else torch.nn.init.zeros_(p)
    def step(i):
        m.zero_grad(set_to_none=True)
        m(*batches[i % 3])[0].sum().backward()
    for i in range(5): step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10): step(i)
    e1.record(); torch.cuda.synchronize()
    print("%s: %.2f ms per step, %.0f molecules/s" % (rnn, e0.elapsed_time(e1) / 10, B * 10 / (e0.elapsed_time(e1) / 1e3)))
