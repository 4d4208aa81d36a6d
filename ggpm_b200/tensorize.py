"""Index builder front end: flat int32 arrays -> the reference's padded batch tensors, built
by the C++ host code in csrc/tensorize.cu behind the C ABI (ggpm_tensorize_graph / _fetch / _cgraph).

Replaces MolGraph.tensorize_graph (ggpm/mol_graph.py:238-281) and the cgraph assembly of MolGraph.tensorize
(:211-222).  The rdkit decomposition that produces the graphs stays out of scope; callers hand over

    node_off   [n_mols+1]          prefix sums of nodes per molecule
    node_label [nodes] or [nodes, 2]
    edge_off   [n_mols+1]          prefix sums of directed edges per molecule
    edges      [n_edges, 4]        (u, v, label, position), u / v local to the molecule, insertion order

and get (fnode, fmess, agraph, bgraph, scope) exactly as the reference lays them out.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence, Tuple

import numpy as np

from . import _lib


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def tensorize_graph_arrays(node_off, node_label, edge_off, edges, with_csr: bool = False):
    node_off, edge_off, edges = _i32(node_off), _i32(edge_off), _i32(edges).reshape(-1, 4)
    node_label = _i32(node_label)
    flat_labels = node_label.ndim == 1
    labels2d = node_label.reshape(node_label.shape[0], -1)
    n_mols = node_off.shape[0] - 1
    if edge_off.shape[0] != n_mols + 1:
        raise ValueError("node_off and edge_off disagree on the number of molecules")
    if labels2d.shape[0] != int(node_off[-1]) or edges.shape[0] != int(edge_off[-1]):
        raise ValueError("offsets do not match the array lengths")
    sizes = np.zeros(6, dtype=np.int32)
    _lib.call_host("ggpm_tensorize_graph", n_mols, _p(node_off), _p(labels2d), labels2d.shape[1], _p(edge_off), _p(edges),
                   _p(sizes))
    n1, e1, wa, wb, nnz_a, nnz_b = (int(v) for v in sizes)
    fnode = np.empty((n1, labels2d.shape[1]), dtype=np.int32)
    fmess = np.empty((e1, 4), dtype=np.int32)
    agraph = np.empty((n1, wa), dtype=np.int32)
    bgraph = np.empty((e1, wb), dtype=np.int32)
    scope = np.empty((n_mols, 2), dtype=np.int32)
    csr = None
    if with_csr:
        csr = (np.empty(n1 + 1, np.int32), np.empty(max(nnz_a, 1), np.int32), np.empty(e1 + 1, np.int32),
               np.empty(max(nnz_b, 1), np.int32))
    _lib.call_host("ggpm_tensorize_fetch", _p(fnode), _p(fmess), _p(agraph), _p(bgraph), _p(scope),
                   *([_p(c) for c in csr] if csr else [None] * 4))
    if flat_labels:
        fnode = fnode.reshape(-1)
    out = (fnode, fmess, agraph, bgraph, [(int(s), int(l)) for s, l in scope])
    if with_csr:
        a_ptr, a_col, b_ptr, b_col = csr
        return out, (a_ptr, a_col[:nnz_a], b_ptr, b_col[:nnz_b])
    return out


def cgraph_arrays(cluster_off, cluster_atoms, motif_mol, tree_scope, graph_scope) -> np.ndarray:
    cluster_off, cluster_atoms, motif_mol = _i32(cluster_off), _i32(cluster_atoms), _i32(motif_mol)
    ts, gs = _i32(tree_scope).reshape(-1, 2), _i32(graph_scope).reshape(-1, 2)
    n_motifs, n_mols = motif_mol.shape[0], ts.shape[0]
    width = np.zeros(1, dtype=np.int32)
    args = (n_mols, n_motifs, _p(cluster_off), _p(cluster_atoms), _p(motif_mol), _p(ts), _p(gs), _p(width))
    _lib.call_host("ggpm_tensorize_cgraph", *args, None)
    cgraph = np.empty((n_motifs + 1, int(width[0])), dtype=np.int32)
    _lib.call_host("ggpm_tensorize_cgraph", *args, _p(cgraph))
    return cgraph


def flatten_edges(node_labels: Sequence[Sequence], edge_lists: Sequence[Sequence[Tuple[int, int, object]]]):
    """lists per molecule (what synth.SynthMol holds; label = int or (int, int)) -> the flat arrays above"""
    node_off = np.zeros(len(node_labels) + 1, dtype=np.int32)
    edge_off = np.zeros(len(edge_lists) + 1, dtype=np.int32)
    np.cumsum([len(l) for l in node_labels], out=node_off[1:])
    np.cumsum([len(e) for e in edge_lists], out=edge_off[1:])
    labels = np.asarray([lab for l in node_labels for lab in l], dtype=np.int32)
    rows: List[Tuple[int, int, int, int]] = []
    for edges in edge_lists:
        for u, v, lab in edges:
            rows.append((u, v, int(lab[0]), int(lab[1])) if isinstance(lab, tuple) else (u, v, int(lab), 0))
    return node_off, labels, edge_off, np.asarray(rows, dtype=np.int32).reshape(-1, 4)


def tensorize(batch):
    """(tree_tensors, graph_tensors) of a list of synth.SynthMol-like molecules (attributes motif_labels, tree_edges,
    atom_labels, bonds, clusters) in the layout of MolGraph.tensorize (ggpm/mol_graph.py:199-236)."""
    tree = tensorize_graph_arrays(*flatten_edges([m.motif_labels for m in batch], [m.tree_edges for m in batch]))
    graph = tensorize_graph_arrays(*flatten_edges([m.atom_labels for m in batch], [m.bonds for m in batch]))
    cluster_off = np.zeros(sum(len(m.clusters) for m in batch) + 1, dtype=np.int32)
    np.cumsum([len(c) for m in batch for c in m.clusters], out=cluster_off[1:])
    atoms = np.asarray([a for m in batch for c in m.clusters for a in c], dtype=np.int32)
    motif_mol = np.asarray([b for b, m in enumerate(batch) for _ in m.clusters], dtype=np.int32)
    cgraph = cgraph_arrays(cluster_off, atoms, motif_mol, tree[4], graph[4])
    return tree[:4] + (cgraph, tree[4]), graph
