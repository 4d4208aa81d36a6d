"""Diagnostic: hidden 500, depth 20 — every GEMM engine against the fp64 oracle, and batch vs halves per engine."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import ggpm_b200
from ggpm_b200 import _lib, ingest, synth, tensorize
from oracle import encoder_oracle as eo
import _checks as C

def model_for(H, depth, seed):
    sd = eo.random_state("hier", "LSTM", H, H, seed)
    m = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, depth, depth, 0.0)
    m.load_state_dict(sd, strict=True)
    return m.cuda(), sd

def run(model, mols):
    tree, graph = ingest.pack(tensorize.tensorize(mols)).to_device("cuda")
    model.zero_grad(set_to_none=True)
    outs = model(tree, graph); outs[0].sum().backward(); torch.cuda.synchronize()
    return [o.detach() for o in outs], {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}

CASES = [tuple(int(v) for v in a.split(',')) for a in sys.argv[1:]] or [(500, 24, 20), (500, 24, 5), (250, 24, 20)]
for H, B, depth in CASES:
    mols = synth.random_batch(int(os.environ.get('DIAG_SEED', '99')), B)
    model, sd = model_for(H, depth, 1)
    tree_np, graph_np = synth.tensorize(mols)
    as_long = lambda ts: [torch.from_numpy(np.asarray(t)).long() if not isinstance(t, list) else t for t in ts]
    torch.set_num_threads(os.cpu_count())
    res = {}
    for dt in (torch.float64, torch.float32):
        sdo = {k: v.to(dt).requires_grad_(True) for k, v in sd.items()}
        ref = eo.hier_encoder(sdo, "LSTM", depth, depth, as_long(tree_np), as_long(graph_np))
        ref[0].sum().backward()
        res[dt] = ([r.detach() for r in ref], {k: v.grad for k, v in sdo.items()})
    r64, r32 = res[torch.float64], res[torch.float32]
    print("H=%d B=%d depth=%d  oracle fp32 vs fp64: out %s  grad max %.2e" % (H, B, depth,
          ["%.1e" % C.rel(a, b) for a, b in zip(r32[0], r64[0])],
          max(C.rel(r32[1][k], g) for k, g in r64[1].items() if g is not None and float(g.abs().max()) > 0)))
    for eng in (1, 2, 0):
        _lib.set_gemm_engine(eng)
        outs, grads = run(model, mols)
        ha, _ = run(model, mols[:B // 2]); hb, _ = run(model, mols[B // 2:])
        errs = sorted(((C.rel(grads[k], g), k) for k, g in r64[1].items() if g is not None and float(g.abs().max()) > 0), reverse=True)
        worst_g = errs[0][0]
        print("      worst grads:", ["%s %.1e" % (k, e) for e, k in errs[:4]])
        halves = [float((torch.cat([ha[i][1 if i else 0:], hb[i][1 if i else 0:]]) - outs[i][1 if i else 0:]).abs().max() / outs[i].abs().max()) for i in range(4)]
        print("   engine %d vs fp64: out %s grad max %.2e | batch vs halves %s" % (eng, ["%.1e" % C.rel(a, b) for a, b in zip(outs, r64[0])], worst_g, ["%.1e" % h for h in halves]))
    _lib.set_gemm_engine(0)
