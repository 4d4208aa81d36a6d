"""TEST INFRASTRUCTURE ONLY — index-construction oracle.

Restates, on real networkx DiGraphs and in the reference's dictionary-based formulation,
  * ``MolGraph.tensorize_graph``      ggpm/mol_graph.py:238-281
  * ``create_pad_tensor``             ggpm/nnutils.py:105-110
  * the ``cgraph`` assembly           ggpm/mol_graph.py:214-222
and defines the CSR re-encoding of a padded index tensor (every non-zero entry, in slot order) that
the CUDA path must reproduce bit-exactly, plus its transpose (for each source row, the destination
rows that gather it, ascending).

Pinned against the live reference's ``tensorize_graph`` by oracle/make_golden.py and
tests/test_tensorize.py (when /root/reference exists) and against tests/golden/*.npz everywhere.
"""
from __future__ import annotations

import networkx as nx
import numpy as np


def to_networkx(node_labels, edges) -> nx.DiGraph:
    """Nodes then edges inserted in the given order, attributes named as the reference expects."""
    g = nx.DiGraph()
    for i, lab in enumerate(node_labels):
        g.add_node(i, label=lab)
    for u, v, lab in edges:
        g.add_edge(u, v, label=lab)
    return g


def pad_rows(rows):
    width = max(len(r) for r in rows) + 1
    return np.asarray([list(r) + [0] * (width - len(r)) for r in rows], dtype=np.int32)


def tensorize_graph_oracle(graphs, vocab):
    """graphs: list of nx.DiGraph with node attr 'label' and edge attr 'label'."""
    fnode, fmess = [None], [(0, 0, 0, 0)]
    into_node, into_mess = [[]], [[]]
    scope, mess_id = [], {}
    for g in graphs:
        first = len(fnode)
        scope.append((first, g.number_of_nodes()))
        g = nx.convert_node_labels_to_integers(g, first_label=first)
        fnode += [None] * g.number_of_nodes()
        for v, lab in g.nodes(data="label"):
            fnode[v] = vocab[lab]
            into_node.append([])
        for u, v, lab in g.edges(data="label"):
            fmess.append((u, v) + (tuple(lab) if isinstance(lab, tuple) else (lab, 0)))
            mess_id[(u, v)] = len(mess_id) + 1
            into_node[v].append(mess_id[(u, v)])
            into_mess.append([])
        for u, v in g.edges:
            into_mess[mess_id[(u, v)]] = [mess_id[(w, u)] for w in g.predecessors(u) if w != v]
    fnode[0] = fnode[1]
    return (np.asarray(fnode, dtype=np.int32), np.asarray(fmess, dtype=np.int32),
            pad_rows(into_node), pad_rows(into_mess), scope)


def cgraph_oracle(clusters_per_mol, tree_scope, graph_scope, n_tree_rows):
    width = max(len(c) for cl in clusters_per_mol for c in cl)
    out = np.zeros((n_tree_rows, width), dtype=np.int32)
    for b, cl in enumerate(clusters_per_mol):
        for i, c in enumerate(cl):
            out[tree_scope[b][0] + i, :len(c)] = [a + graph_scope[b][0] for a in c]
    return out


def csr_from_padded_oracle(index: np.ndarray):
    """(row_ptr [M+1], col [nnz]) of the non-zero entries of a padded [M, K] index tensor, slot order."""
    index = np.asarray(index)
    keep = index != 0
    row_ptr = np.zeros(index.shape[0] + 1, dtype=np.int32)
    row_ptr[1:] = np.cumsum(keep.sum(1))
    return row_ptr, index[keep].astype(np.int32)


def csr_transpose_oracle(row_ptr: np.ndarray, col: np.ndarray, n_src_rows: int):
    """For every source row n: the destination rows e with n in row e's list, ascending (duplicates
    kept), and for each such pair the position of the pair in the forward ``col`` array."""
    dest = np.repeat(np.arange(len(row_ptr) - 1, dtype=np.int32), np.diff(row_ptr))
    order = np.lexsort((np.arange(len(col)), dest, col))          # by source, then dest, then slot
    t_ptr = np.zeros(n_src_rows + 1, dtype=np.int32)
    t_ptr[1:] = np.cumsum(np.bincount(col, minlength=n_src_rows))
    return t_ptr, dest[order].astype(np.int32), order.astype(np.int32)
