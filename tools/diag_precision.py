"""Where does the gradient error of a sweep-size batch come from?  (B200 only; test infrastructure: imports tests/ and oracle/)

    python tools/diag_precision.py --hidden 500 --batch 1024 --seed 31024

Runs the product under each GEMM engine (1 = fp32 CUDA cores, 2 = tcgen05 3xTF32, 0 = auto: f16x3 for the LSTM) on ONE
seeded batch, reads the ReLU pattern it took, evaluates the fp64 oracle on the GPU on THAT pattern (so kink effects are
out of the picture) and the reference arithmetic in fp32 on the same pattern, and prints max|g - ref| / max|ref| per
parameter.  The fp32 column is the noise floor the reference's own arithmetic has on this batch.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("GGPM_ALLOW_TEST_BACKEND", "1")

import torch                                                  # noqa: E402

import _checks as C                                           # noqa: E402
import test_gpu_parity_full as T                              # noqa: E402
from ggpm_b200 import _lib, synth                             # noqa: E402
from oracle import encoder_oracle as eo                       # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=500)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--depth", type=int, default=20)
    ap.add_argument("--seed", type=int, default=31024)
    ap.add_argument("--engines", default="0,2,1")
    a = ap.parse_args()
    kind, rnn, H, depth = "hier", "LSTM", a.hidden, a.depth
    mols = synth.random_batch(a.seed, a.batch)
    sd = eo.random_state(kind, rnn, H, H, a.seed % 97)
    tree_np, graph_np = synth.tensorize(mols)
    print("batch %d, hidden %d, depth %d: %d atom-layer message rows" % (a.batch, H, depth, graph_np[1].shape[0]))
    cols, names = {}, None
    for eng in [int(e) for e in a.engines.split(",")]:
        _lib.set_gemm_engine(eng)
        outs, grads, masks = T._product(kind, rnn, H, depth, depth, sd, mols, T._root_sum)
        _lib.set_gemm_engine(0)
        outs, grads = [o.cpu() for o in outs], {k: None if g is None else g.cpu() for k, g in grads.items()}
        torch.cuda.empty_cache()
        relu = lambda dt: {k: v.to(dt) for k, v in masks.items()}
        o64, g64, _ = T._oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float64, T._root_sum, relu=relu(torch.float64), device="cuda")
        o32, g32, _ = T._oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float32, T._root_sum, relu=relu(torch.float32), device="cuda")
        names = [k for k, g in g64.items() if g is not None and float(g.abs().max()) > 0]
        cols["engine %d" % eng] = {k: C.rel(grads[k], g64[k]) for k in names}
        cols["ref fp32 (pattern of engine %d)" % eng] = {k: C.rel(g32[k], g64[k]) for k in names}
        cols["engine %d" % eng]["out:hroot"] = C.rel(outs[0], o64[0])
        cols["ref fp32 (pattern of engine %d)" % eng]["out:hroot"] = C.rel(o32[0], o64[0])
    heads = list(cols)
    print("%-36s" % "tensor" + "".join("%12s" % h.replace("ref fp32 (pattern of engine ", "fp32@e").replace(")", "") for h in heads))
    for k in ["out:hroot"] + names:
        print("%-36s" % k + "".join("%12.2e" % cols[h][k] for h in heads))
    print("%-36s" % "max over gradients" + "".join("%12.2e" % max(v for k, v in cols[h].items() if k != "out:hroot") for h in heads))


if __name__ == "__main__":
    main()
