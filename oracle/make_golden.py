"""TEST INFRASTRUCTURE ONLY — generate tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (where /root/reference exists):

    python -m oracle.make_golden

For every case below it
  1. draws a seeded synthetic batch (ggpm_b200.synth), turns it into networkx DiGraphs and runs the
     reference's own ``MolGraph.tensorize_graph`` (ggpm/mol_graph.py:238-281) on them; asserts the
     product tensoriser and the oracle tensoriser reproduce those tensors bit-for-bit;
  2. builds the reference ``HierMPNEncoder`` / ``MotifEncoder`` (ggpm/encoder.py:41-157, 252-341) in
     fp64, strictly loads a seeded state_dict (fp32-representable values), runs forward and backward
     (random cotangents on every output, and separately ``hroot.sum()``);
  3. asserts oracle/encoder_oracle.py agrees with the reference to <= 1e-12 (outputs and every
     parameter gradient) — this is the oracle's parity pin;
  4. writes inputs, outputs and gradients to tests/golden/<case>.npz.

The weights are not stored: they are regenerated from the seed by
``oracle.encoder_oracle.random_state`` (torch CPU generator, same image on every box).
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ggpm_b200 import synth                                    # noqa: E402
from oracle import encoder_oracle as eo                        # noqa: E402
from oracle import tensorize_oracle as to                      # noqa: E402
from oracle.ref_loader import FakePairVocab, IdAtomVocab, load_reference  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# name, kind, rnn, hidden, depthT, depthG, batch, data seed, weight seed, (n_motif, n_attach), full grads?
CASES = [
    dict(name="hier_lstm_small", kind="hier", rnn="LSTM", hidden=32, depthT=4, depthG=5, batch=6,
         data_seed=11, weight_seed=101, vocab=(40, 60), mean_motifs=5.0, full=True),
    dict(name="hier_gru_small", kind="hier", rnn="GRU", hidden=32, depthT=4, depthG=5, batch=6,
         data_seed=12, weight_seed=102, vocab=(40, 60), mean_motifs=5.0, full=True),
    dict(name="motif_lstm_small", kind="motif", rnn="LSTM", hidden=24, depthT=6, depthG=0, batch=5,
         data_seed=13, weight_seed=103, vocab=(40, 60), mean_motifs=6.0, full=True),
    dict(name="motif_gru_small", kind="motif", rnn="GRU", hidden=24, depthT=6, depthG=0, batch=5,
         data_seed=14, weight_seed=104, vocab=(40, 60), mean_motifs=6.0, full=True),
    dict(name="hier_lstm_deep", kind="hier", rnn="LSTM", hidden=40, depthT=20, depthG=20, batch=8,
         data_seed=15, weight_seed=105, vocab=(40, 60), mean_motifs=8.0, full=True),
    dict(name="hier_gru_deep", kind="hier", rnn="GRU", hidden=40, depthT=20, depthG=20, batch=8,
         data_seed=16, weight_seed=106, vocab=(40, 60), mean_motifs=8.0, full=True),
    # BASELINE config #1 stand-in: configs/configs.json hyper-parameters, 20 synthetic molecules
    dict(name="hier_lstm_cfg1", kind="hier", rnn="LSTM", hidden=250, depthT=20, depthG=20, batch=20,
         data_seed=1, weight_seed=0, vocab=(721, 6215), mean_motifs=12.0, full=False),
    # edge cases of the motif tree: single-motif molecules (no tree edge) mixed into a batch, and batches
    # whose tree message table is the padding row alone (Et = 0, bgraph / agraph of width 1)
    dict(name="hier_lstm_lone_motifs", kind="hier", rnn="LSTM", hidden=32, depthT=4, depthG=5, batch=7,
         data_seed=17, weight_seed=107, vocab=(40, 60), mean_motifs=1.6, sd_motifs=0.8, min_motifs=1, full=True),
    dict(name="hier_lstm_no_tree_edges", kind="hier", rnn="LSTM", hidden=32, depthT=4, depthG=5, batch=4,
         data_seed=18, weight_seed=108, vocab=(40, 60), mean_motifs=1.0, sd_motifs=0.0, min_motifs=1, full=True),
    dict(name="hier_gru_no_tree_edges", kind="hier", rnn="GRU", hidden=32, depthT=4, depthG=5, batch=4,
         data_seed=19, weight_seed=109, vocab=(40, 60), mean_motifs=1.0, sd_motifs=0.0, min_motifs=1, full=True),
    dict(name="motif_lstm_no_tree_edges", kind="motif", rnn="LSTM", hidden=24, depthT=6, depthG=0, batch=3,
         data_seed=20, weight_seed=110, vocab=(40, 60), mean_motifs=1.0, sd_motifs=0.0, min_motifs=1, full=True),
]
GRAD_STRIDE = 97      # big-case gradients are stored as flat[::GRAD_STRIDE] plus their L2 norm


def build_inputs(case):
    nm, na = case["vocab"]
    mols = synth.random_batch(case["data_seed"], case["batch"], n_motif_vocab=nm, n_attach_vocab=na,
                              mean_motifs=case["mean_motifs"],
                              sd_motifs=case.get("sd_motifs", case["mean_motifs"] / 4.0),
                              min_motifs=case.get("min_motifs", 2))
    return mols, synth.tensorize(mols)


def reference_tensorize(ref, mols, vocab):
    trees = [to.to_networkx(m.motif_labels, m.tree_edges) for m in mols]
    graphs = [to.to_networkx(m.atom_labels, m.bonds) for m in mols]
    tree, _ = ref.mol_graph.MolGraph.tensorize_graph(trees, vocab)
    graph, _ = ref.mol_graph.MolGraph.tensorize_graph(graphs, IdAtomVocab())
    return tree, graph, trees, graphs


def as_long(tensors):
    return [torch.as_tensor(np.asarray(t)).long() if not isinstance(t, list) else t for t in tensors]


def rel_err(a, b):
    denom = max(float(b.abs().max()), 1e-300)
    return float((a - b).abs().max()) / denom


def run_case(ref, case):
    nm, na = case["vocab"]
    vocab = FakePairVocab(nm, na)
    mols, (tree_np, graph_np) = build_inputs(case)

    # --- index construction: reference vs product tensoriser vs oracle tensoriser ---------------
    r_tree, r_graph, nx_trees, nx_graphs = reference_tensorize(ref, mols, vocab)
    o_tree = to.tensorize_graph_oracle(nx_trees, vocab)
    o_graph = to.tensorize_graph_oracle(nx_graphs, IdAtomVocab())
    for k in range(4):
        assert np.array_equal(r_tree[k].numpy(), tree_np[k]), ("tree", k)
        assert np.array_equal(r_graph[k].numpy(), graph_np[k]), ("graph", k)
        assert np.array_equal(r_tree[k].numpy(), o_tree[k]), ("otree", k)
        assert np.array_equal(r_graph[k].numpy(), o_graph[k]), ("ograph", k)
    assert r_tree[4] == tree_np[5] == o_tree[4] and r_graph[4] == graph_np[4] == o_graph[4]
    cg = to.cgraph_oracle([m.clusters for m in mols], r_tree[4], r_graph[4], r_tree[0].shape[0])
    assert np.array_equal(cg, tree_np[4])

    tree_t, graph_t = as_long(tree_np), as_long(graph_np)

    # --- reference model in fp64 ----------------------------------------------------------------
    sd32 = eo.random_state(case["kind"], case["rnn"], case["hidden"], case["hidden"], case["weight_seed"],
                           n_motif=nm, n_attach=na)
    sd64 = {k: v.double() for k, v in sd32.items()}
    torch.set_default_dtype(torch.float64)
    try:
        cls = ref.encoder.HierMPNEncoder if case["kind"] == "hier" else ref.encoder.MotifEncoder
        model = cls(vocab, IdAtomVocab(), case["rnn"], case["hidden"], case["hidden"], case["depthT"],
                    case["depthG"], 0.0)
        assert list(model.state_dict().keys()) == list(sd64.keys()), "state_dict key order differs"
        model.load_state_dict(sd64, strict=True)
        model.double()

        def ref_forward():
            return model(tree_t, graph_t) if case["kind"] == "hier" else model(tree_t)

        outs = ref_forward()
        gen = torch.Generator().manual_seed(case["weight_seed"] + 7)
        cots = [torch.randn(o.shape, generator=gen, dtype=torch.float64) for o in outs]
        model.zero_grad()
        sum((o * c).sum() for o, c in zip(outs, cots)).backward()
        grads_all = {k: (p.grad.clone() if p.grad is not None else torch.zeros_like(p))
                     for k, p in model.named_parameters()}
        model.zero_grad()
        ref_forward()[0].sum().backward()
        grads_root = {k: (p.grad.clone() if p.grad is not None else None) for k, p in model.named_parameters()}
    finally:
        torch.set_default_dtype(torch.float32)

    # --- oracle must agree with the reference (parity pin) --------------------------------------
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd64.items()}
    if case["kind"] == "hier":
        o_outs = eo.hier_encoder(sdo, case["rnn"], case["depthT"], case["depthG"], tree_t, graph_t)
    else:
        o_outs = eo.motif_encoder(sdo, case["rnn"], case["depthT"], tree_t)
    worst = max(rel_err(a, b) for a, b in zip(o_outs, outs))
    sum((o * c).sum() for o, c in zip(o_outs, cots)).backward()
    for k in sdo:
        g = sdo[k].grad if sdo[k].grad is not None else torch.zeros_like(sdo[k])
        if float(grads_all[k].abs().max()) > 0:
            worst = max(worst, rel_err(g, grads_all[k]))
        else:
            assert float(g.abs().max()) == 0.0, k
    assert worst <= 1e-12, (case["name"], worst)

    # --- reference in fp32 (records the fp32-vs-fp64 headroom of the reference itself) ----------
    model32 = cls(vocab, IdAtomVocab(), case["rnn"], case["hidden"], case["hidden"], case["depthT"],
                  case["depthG"], 0.0)
    model32.load_state_dict(sd32, strict=True)
    with torch.no_grad():
        outs32 = model32(tree_t, graph_t) if case["kind"] == "hier" else model32(tree_t)
    fp32_gap = max(rel_err(a.double(), b) for a, b in zip(outs32, outs))

    # --- write ----------------------------------------------------------------------------------
    names = ("hroot", "hnode", "hinter", "hatom") if case["kind"] == "hier" else ("hroot", "hnode")
    blob = {"meta": np.frombuffer(json.dumps(dict(case, fp32_gap=fp32_gap, oracle_vs_ref=worst,
                                                  grad_stride=GRAD_STRIDE)).encode(), dtype=np.uint8)}
    for i, nme in enumerate(("fnode", "fmess", "agraph", "bgraph", "cgraph")):
        blob["tree_" + nme] = tree_np[i]
    blob["tree_scope"] = np.asarray(tree_np[5], dtype=np.int32)
    for i, nme in enumerate(("fnode", "fmess", "agraph", "bgraph")):
        blob["graph_" + nme] = graph_np[i]
    blob["graph_scope"] = np.asarray(graph_np[4], dtype=np.int32)
    for nme, o, c in zip(names, outs, cots):
        blob["out_" + nme] = o.detach().numpy() if case["full"] else o.detach().numpy().astype(np.float32)
        if case["full"]:
            blob["cot_" + nme] = c.numpy()
    for k in grads_all:
        if case["full"]:
            blob["gall_" + k] = grads_all[k].numpy()
        gr = grads_root[k]
        if gr is None:
            continue                      # parameter receives no gradient from hroot (tree_encoder.W_o)
        if case["full"]:
            blob["groot_" + k] = gr.numpy()
        else:
            blob["groot_" + k] = gr.reshape(-1)[::GRAD_STRIDE].numpy().copy()
            blob["gnorm_" + k] = np.asarray(float(gr.norm()))
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, case["name"] + ".npz")
    np.savez_compressed(path, **blob)
    print("%-18s oracle-vs-ref %.2e  ref fp32-vs-fp64 %.2e  -> %s (%.0f KB)" %
          (case["name"], worst, fp32_gap, os.path.relpath(path, ROOT), os.path.getsize(path) / 1024))


def main():
    ref = load_reference()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    only = set(sys.argv[1:])          # optional: names of the cases to (re)generate
    for case in CASES:
        if not only or case["name"] in only:
            run_case(ref, case)


if __name__ == "__main__":
    main()
