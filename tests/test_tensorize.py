"""CPU: the product tensoriser (ggpm_b200.synth) is bit-exact against the index-construction oracle,
the golden tensors made by the reference's tensorize_graph, and (when present) the live reference."""
import numpy as np
import pytest

from conftest import golden_names, load_golden
from ggpm_b200 import synth
from oracle import tensorize_oracle as to
from oracle.ref_loader import FakePairVocab, IdAtomVocab, load_reference, reference_available


def _regen(meta):
    nm, na = meta["vocab"]
    mols = synth.random_batch(meta["data_seed"], meta["batch"], n_motif_vocab=nm, n_attach_vocab=na,
                              mean_motifs=meta["mean_motifs"],
                              sd_motifs=meta.get("sd_motifs", meta["mean_motifs"] / 4.0),
                              min_motifs=meta.get("min_motifs", 2))
    return mols, synth.tensorize(mols)


@pytest.mark.parametrize("name", golden_names())
def test_tensorize_matches_golden(name):
    meta, z, _, _ = load_golden(name)
    mols, (tree, graph) = _regen(meta)
    for i, k in enumerate(("fnode", "fmess", "agraph", "bgraph", "cgraph")):
        assert np.array_equal(tree[i], z["tree_" + k]), k
        assert tree[i].dtype == np.int32
    for i, k in enumerate(("fnode", "fmess", "agraph", "bgraph")):
        assert np.array_equal(graph[i], z["graph_" + k]), k
    assert [list(s) for s in tree[5]] == z["tree_scope"].tolist()
    assert [list(s) for s in graph[4]] == z["graph_scope"].tolist()


@pytest.mark.parametrize("seed,batch", [(0, 1), (3, 7), (9, 33)])
def test_tensorize_matches_oracle_on_networkx(seed, batch):
    mols = synth.random_batch(seed, batch)
    tree, graph = synth.tensorize(mols)
    o_tree = to.tensorize_graph_oracle([to.to_networkx(m.motif_labels, m.tree_edges) for m in mols],
                                       FakePairVocab())
    o_graph = to.tensorize_graph_oracle([to.to_networkx(m.atom_labels, m.bonds) for m in mols], IdAtomVocab())
    for k in range(4):
        assert np.array_equal(tree[k], o_tree[k])
        assert np.array_equal(graph[k], o_graph[k])
    assert tree[5] == o_tree[4] and graph[4] == o_graph[4]
    cg = to.cgraph_oracle([m.clusters for m in mols], o_tree[4], o_graph[4], o_tree[0].shape[0])
    assert np.array_equal(cg, tree[4])


@pytest.mark.skipif(not reference_available(), reason="/root/reference not present on this box")
def test_tensorize_matches_live_reference():
    ref = load_reference()
    mols = synth.random_batch(21, 9)
    tree, graph = synth.tensorize(mols)
    r_tree, _ = ref.mol_graph.MolGraph.tensorize_graph(
        [to.to_networkx(m.motif_labels, m.tree_edges) for m in mols], FakePairVocab())
    r_graph, _ = ref.mol_graph.MolGraph.tensorize_graph(
        [to.to_networkx(m.atom_labels, m.bonds) for m in mols], IdAtomVocab())
    for k in range(4):
        assert np.array_equal(tree[k], r_tree[k].numpy())
        assert np.array_equal(graph[k], r_graph[k].numpy())


def test_layout_invariants():
    mols = synth.random_batch(5, 12)
    tree, graph = synth.tensorize(mols)
    for fnode, fmess, agraph, bgraph in (tree[:4], graph[:4]):
        assert (agraph[:, -1] == 0).all() and (bgraph[:, -1] == 0).all()      # >= 1 trailing zero
        assert (agraph[0] == 0).all() and (bgraph[0] == 0).all() and (fmess[0] == 0).all()
        assert np.array_equal(fnode[0], fnode[1])
        for idx in (agraph, bgraph):                                           # zeros only trailing
            nz = idx != 0
            assert (nz[:, :-1] >= nz[:, 1:]).all()
        # bgraph[e=(u->v)] lists messages (w->u) with w != v
        e = len(fmess) - 1
        u, v = fmess[e, 0], fmess[e, 1]
        for p in bgraph[e][bgraph[e] != 0]:
            assert fmess[p, 1] == u and fmess[p, 0] != v


def test_csr_oracle_roundtrip():
    mols = synth.random_batch(2, 4)
    _, graph = synth.tensorize(mols)
    bg = graph[3]
    rp, col = to.csr_from_padded_oracle(bg)
    assert rp[-1] == np.count_nonzero(bg)
    for r in (0, 1, len(bg) - 1):
        assert list(col[rp[r]:rp[r + 1]]) == [x for x in bg[r] if x != 0]
    tp, tdest, tpos = to.csr_transpose_oracle(rp, col, len(bg))
    assert tp[-1] == len(col)
    for n in range(1, len(bg), 17):
        assert sorted(tdest[tp[n]:tp[n + 1]]) == [e for e in range(len(bg)) if n in bg[e]]
        assert (col[tpos[tp[n]:tp[n + 1]]] == n).all()
