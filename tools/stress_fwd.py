"""Stress: repeat forward (+backward) of the same batch many times; every repeat must reproduce the first outputs to
rounding.  Usage: python tools/stress_fwd.py H B iters [bwd]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggpm_b200
from ggpm_b200 import ingest, synth, tensorize
from oracle import encoder_oracle as eo
H, B, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
bwd = len(sys.argv) > 4
sd = eo.random_state("hier", "LSTM", H, H, 1)
model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, 20, 20, 0.0)
model.load_state_dict(sd, strict=True); model = model.cuda()
batches = [synth.random_batch(777 + H, B), synth.random_batch(778 + H, B // 2), synth.random_batch(779 + H, B // 3)]
packs = [ingest.pack(tensorize.tensorize(m)) for m in batches]
ref = [None] * len(packs)
bad = 0
t0 = time.time()
for it in range(iters):
    k = it % len(packs)
    tree, graph = packs[k].to_device("cuda")
    if bwd:
        model.zero_grad(set_to_none=True)
        outs = model(tree, graph); outs[0].sum().backward()
        res = [o.detach() for o in outs] + [model.graph_encoder.rnn.W_i[0].weight.grad, model.tree_encoder.rnn.W_f[0].weight.grad]
    else:
        with torch.no_grad():
            res = list(model(tree, graph))
    res = [r.clone() for r in res]
    if ref[k] is None:
        ref[k] = res
        continue
    errs = [float((a - b).abs().max() / b.abs().max()) for a, b in zip(res, ref[k])]
    if max(errs) > 2e-5:
        bad += 1
        print("iter %d batch %d: %s" % (it, k, ["%.1e" % e for e in errs]), flush=True)
print("H=%d B=%d iters=%d bwd=%s PDL=%s: %d mismatching repeats, %.1f s" % (H, B, iters, bwd, os.environ.get("GGPM_PDL", "1"), bad, time.time() - t0))
