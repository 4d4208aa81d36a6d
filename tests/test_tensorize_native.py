"""CPU: the C++ index builder behind the C ABI (csrc/tensorize.cu, ggpm_b200/tensorize.py)
against the golden tensors made by the reference's own MolGraph.tensorize_graph, the index oracle on networkx graphs,
and the Python builder of the synthetic generator — bit for bit."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import GOLDEN_DIR, golden_names                                          # noqa: E402
from ggpm_b200 import ingest, synth, tensorize                                         # noqa: E402
from oracle import tensorize_oracle as to                                              # noqa: E402


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if isinstance(x, list):
            assert [tuple(s) for s in x] == [tuple(s) for s in y]
        else:
            x, y = np.asarray(x), np.asarray(y)
            assert x.shape == y.shape and x.dtype == np.int32 and np.array_equal(x, y)


@pytest.mark.parametrize("seed,n", [(1, 1), (2, 3), (5, 20), (9, 256)])
def test_native_builder_equals_python_builder(seed, n):
    mols = synth.random_batch(seed, n)
    ref_tree, ref_graph = synth.tensorize(mols)
    tree, graph = tensorize.tensorize(mols)
    _same(tree, ref_tree)
    _same(graph, ref_graph)


@pytest.mark.parametrize("seed,n", [(3, 4), (4, 17)])
def test_native_builder_equals_oracle_on_networkx(seed, n):
    """the oracle walks nx.DiGraph objects exactly as the reference does (ggpm/mol_graph.py:238-281)"""
    mols = synth.random_batch(seed, n)
    for labels, edges in (([m.atom_labels for m in mols], [m.bonds for m in mols]),
                          ([m.motif_labels for m in mols], [m.tree_edges for m in mols])):
        graphs = [to.to_networkx(l, e) for l, e in zip(labels, edges)]
        ofn, ofm, oag, obg, oscope = to.tensorize_graph_oracle(graphs, to.IdentityVocab() if hasattr(to, "IdentityVocab") else _Id())
        fn, fm, ag, bg, scope = tensorize.tensorize_graph_arrays(*tensorize.flatten_edges(labels, edges))
        assert np.array_equal(np.asarray(ofn), fn) and np.array_equal(np.asarray(ofm), fm)
        assert np.array_equal(np.asarray(oag), ag) and np.array_equal(np.asarray(obg), bg)
        assert [tuple(s) for s in oscope] == scope


class _Id:
    def __getitem__(self, k):
        return k


def test_csr_output_matches_padded_output():
    mols = synth.random_batch(8, 12)
    labels, edges = [m.atom_labels for m in mols], [m.bonds for m in mols]
    (fn, fm, ag, bg, scope), (a_ptr, a_col, b_ptr, b_col) = tensorize.tensorize_graph_arrays(
        *tensorize.flatten_edges(labels, edges), with_csr=True)
    for padded, ptr, col in ((ag, a_ptr, a_col), (bg, b_ptr, b_col)):
        rp, cc = ingest.csr_from_padded_host(padded)
        assert np.array_equal(rp, ptr) and np.array_equal(cc, col)
    assert ag.shape[1] == np.diff(a_ptr).max() + 1 and bg.shape[1] == np.diff(b_ptr).max() + 1   # one trailing zero column
    assert fn[0] == fn[1] and not fm[0].any() and not ag[0].any() and not bg[0].any()


def test_edge_cases_and_errors():
    # one molecule, one node, no edges
    fn, fm, ag, bg, scope = tensorize.tensorize_graph_arrays([0, 1], [7], [0, 0], np.zeros((0, 4), np.int32))
    assert fn.tolist() == [7, 7] and fm.shape == (1, 4) and ag.shape == (2, 1) and bg.shape == (1, 1) and scope == [(1, 1)]
    # two-column labels (motif, attachment), an empty molecule in the middle
    fn, fm, ag, bg, scope = tensorize.tensorize_graph_arrays([0, 2, 2, 3], [[1, 2], [3, 4], [5, 6]], [0, 2, 2, 2],
                                                             [[0, 1, 0, 0], [1, 0, 1, 0]])
    assert fn.tolist() == [[1, 2], [1, 2], [3, 4], [5, 6]] and scope == [(1, 2), (3, 0), (3, 1)]
    assert fm.tolist() == [[0, 0, 0, 0], [1, 2, 0, 0], [2, 1, 1, 0]] and ag.tolist() == [[0, 0], [2, 0], [1, 0], [0, 0]]
    assert bg.tolist() == [[0], [0], [0]]                      # the only predecessor of each source is the edge's own target
    with pytest.raises(RuntimeError):                         # node id outside the molecule
        tensorize.tensorize_graph_arrays([0, 2], [1, 1], [0, 1], [[0, 2, 0, 0]])
    with pytest.raises(ValueError):
        tensorize.tensorize_graph_arrays([0, 2], [1, 1], [0, 1, 1], [[0, 1, 0, 0]])


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("hier")][:3])
def test_golden_index_tensors_are_self_consistent_with_the_builder(name):
    """golden tensors came out of the reference's tensorize_graph: rebuilding them from their own fmess must be the
    identity (edge list recovered from fmess in id order = enumeration order)"""
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    for fam in ("graph", "tree"):
        fmess, scope = z[fam + "_fmess"], z[fam + "_scope"]
        fnode = z[fam + "_fnode"]
        node_off = np.concatenate([[0], np.cumsum(scope[:, 1])]).astype(np.int32)
        start_of = {int(s): b for b, (s, l) in enumerate(scope)}
        mol_of_node = np.zeros(fnode.shape[0], dtype=np.int64)
        for b, (s, l) in enumerate(scope):
            mol_of_node[s:s + l] = b
        e_mol = mol_of_node[fmess[1:, 0]]
        edge_off = np.concatenate([[0], np.cumsum(np.bincount(e_mol, minlength=len(scope)))]).astype(np.int32)
        local = fmess[1:].copy()
        local[:, 0] -= scope[e_mol, 0]
        local[:, 1] -= scope[e_mol, 0]
        fn, fm, ag, bg, sc = tensorize.tensorize_graph_arrays(node_off, fnode[1:], edge_off, local)
        assert np.array_equal(fn, fnode) and np.array_equal(fm, fmess)
        assert np.array_equal(ag, z[fam + "_agraph"]) and np.array_equal(bg, z[fam + "_bgraph"])
