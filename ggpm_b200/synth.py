"""Synthetic molecular graphs and the batch tensoriser (host side, numpy only).

Produces the encoder's input format — the padded index tensors that the reference builds with
``MolGraph.tensorize`` / ``MolGraph.tensorize_graph`` (ggpm/mol_graph.py:199-281) and
``create_pad_tensor`` (ggpm/nnutils.py:105-110) — from seeded synthetic molecules, because the
real chemistry (rdkit decomposition, ggpm/mol_graph.py:24-197) is out of scope and rdkit is not
available.  The generator follows SURVEY.md §8(d).

Layout facts reproduced bit-for-bit (verified against the reference on networkx graphs in
tests/test_tensorize.py):
  * node / message ids are 1-based over the whole batch, id 0 is padding;
  * messages are enumerated per molecule in ``G.edges`` order of a networkx DiGraph, i.e. grouped
    by source node in node order, out-edges in insertion order;
  * ``agraph[v]`` = ids of messages into v, ``bgraph[e=(u->v)]`` = ids of messages (w->u), w != v,
    both in ascending id order; rows are zero padded to ``max_len + 1``;
  * ``fnode[0] = fnode[1]``, ``fmess[0] = (0, 0, 0, 0)``;
  * ``cgraph[v]`` = batch-global atom ids of motif v's cluster, zero padded to the largest
    cluster (no extra column), ggpm/mol_graph.py:214-222.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple, Union

import numpy as np

BondLabel = Union[int, Tuple[int, int]]


@dataclass
class SynthMol:
    """One molecule: an atom graph and its motif tree, both as ordered directed edge lists."""
    atom_labels: List[int] = field(default_factory=list)          # index into the atom vocab
    bonds: List[Tuple[int, int, BondLabel]] = field(default_factory=list)   # directed, insertion order
    motif_labels: List[Tuple[int, int]] = field(default_factory=list)       # (motif id, attachment id)
    tree_edges: List[Tuple[int, int, int]] = field(default_factory=list)    # directed, insertion order
    clusters: List[List[int]] = field(default_factory=list)       # atoms of each motif


def _edges_in_nx_order(n_nodes: int, edges: Sequence[Tuple[int, int, object]]):
    """Directed edges in the order ``nx.DiGraph.edges`` yields them: by source in node order,
    then by insertion order of that source's out-edges."""
    out = [[] for _ in range(n_nodes)]
    for u, v, lab in edges:
        out[u].append((u, v, lab))
    return [e for lst in out for e in lst]


def random_molecule(rng: random.Random, n_motif_vocab=721, n_attach_vocab=6215, n_atom_vocab=38,
                    mean_motifs=12.0, sd_motifs=3.0, min_motifs=2) -> SynthMol:
    """Degree-bounded random molecule: rings (5/6/6 atoms, aromatic bonds) and 2-atom bonds glued
    by sharing one atom, exactly one tree edge per glue.  ``min_motifs=1`` admits
    single-motif molecules, whose motif tree has no edge at all (a lone ring in the real data)."""
    mol = SynthMol()
    degree: List[int] = []
    n_motifs = max(min_motifs, int(round(rng.gauss(mean_motifs, sd_motifs))))

    def new_atom():
        mol.atom_labels.append(rng.randrange(n_atom_vocab))
        degree.append(0)
        return len(mol.atom_labels) - 1

    def add_bond(u, v, btype):
        lab: BondLabel = btype
        if rng.random() < 0.2:
            lab = (btype, rng.randint(1, 4))
        mol.bonds.append((u, v, lab))
        mol.bonds.append((v, u, lab))
        degree[u] += 1
        degree[v] += 1

    def add_motif(shared_atom, is_ring):
        if is_ring:
            size = rng.choice((5, 6, 6))
            atoms = [shared_atom if shared_atom is not None else new_atom()]
            atoms += [new_atom() for _ in range(size - 1)]
            for i in range(size):
                add_bond(atoms[i], atoms[(i + 1) % size], 3)
        else:
            a = shared_atom if shared_atom is not None else new_atom()
            atoms = [a, new_atom()]
            add_bond(atoms[0], atoms[1], rng.randrange(3))
        mol.clusters.append(atoms)
        mol.motif_labels.append((rng.randrange(n_motif_vocab), rng.randrange(n_attach_vocab)))
        return len(mol.clusters) - 1

    add_motif(None, rng.random() < 0.45)
    children: List[List[int]] = [[]]
    for _ in range(1, n_motifs):
        is_ring = rng.random() < 0.45
        cap = 2 if is_ring else 3          # a ring adds two bonds to the shared atom, a bond adds one
        for _attempt in range(16):
            pm = rng.randrange(len(mol.clusters))
            cands = [a for a in mol.clusters[pm] if degree[a] <= cap]
            if not cands:
                continue
            cm = add_motif(rng.choice(cands), is_ring)
            children.append([])
            children[pm].append(cm)
            mol.tree_edges.append((pm, cm, 0))
            mol.tree_edges.append((cm, pm, 0))
            break

    # tree edge labels by DFS from motif 0 (ggpm/mol_graph.py:121-134): parent->child 0,
    # child->parent = sibling index + 1 with children visited in sorted order
    label = {}
    stack = [0]
    while stack:
        x = stack.pop()
        for idx, y in enumerate(sorted(children[x])):
            label[(x, y)] = 0
            label[(y, x)] = idx + 1
            stack.append(y)
    mol.tree_edges = [(u, v, label[(u, v)]) for (u, v, _) in mol.tree_edges]
    return mol


def random_batch(seed: int, batch_size: int, **kw) -> List[SynthMol]:
    rng = random.Random(seed)
    return [random_molecule(rng, **kw) for _ in range(batch_size)]


def _pad_rows(rows: List[List[int]], extra: int = 1) -> np.ndarray:
    width = max(len(r) for r in rows) + extra
    out = np.zeros((len(rows), width), dtype=np.int32)
    for i, r in enumerate(rows):
        if r:
            out[i, :len(r)] = r
    return out


def tensorize_graph(node_labels: Sequence[Sequence], edge_lists: Sequence[Sequence[Tuple[int, int, object]]]):
    """Batch-level padded index tensors for one graph family (atom graphs or motif trees).

    node_labels[b][i]  : label of node i of molecule b (int, or (int, int) for motifs)
    edge_lists[b]      : directed (u, v, label) edges of molecule b in insertion order
    Returns (fnode, fmess, agraph, bgraph, scope) as int32 arrays + list of (start, length).
    Restates ggpm/mol_graph.py:238-281.
    """
    fnode: List = [None]
    fmess: List[Tuple[int, int, int, int]] = [(0, 0, 0, 0)]
    agraph: List[List[int]] = [[]]
    bgraph: List[List[int]] = [[]]
    scope: List[Tuple[int, int]] = []
    for labels, edges in zip(node_labels, edge_lists):
        offset = len(fnode)
        n = len(labels)
        scope.append((offset, n))
        fnode.extend(labels)
        agraph.extend([] for _ in range(n))
        first_eid = len(fmess)
        ordered = _edges_in_nx_order(n, edges)
        eid_of = {}
        for u, v, lab in ordered:
            if isinstance(lab, tuple):
                fmess.append((u + offset, v + offset, int(lab[0]), int(lab[1])))
            else:
                fmess.append((u + offset, v + offset, int(lab), 0))
            eid = len(fmess) - 1
            eid_of[(u, v)] = eid
            agraph[v + offset].append(eid)
            bgraph.append([])
        # predecessors of u in the relabelled DiGraph come in edge-enumeration order
        into: List[List[Tuple[int, int]]] = [[] for _ in range(n)]
        for u, v, _ in ordered:
            into[v].append((u, eid_of[(u, v)]))
        for u, v, _ in ordered:
            eid = eid_of[(u, v)]
            bgraph[eid] = [e for (w, e) in into[u] if w != v]
        assert len(fmess) - first_eid == len(ordered)
    fnode[0] = fnode[1]
    fnode_arr = np.asarray(fnode, dtype=np.int32)
    fmess_arr = np.asarray(fmess, dtype=np.int32)
    return fnode_arr, fmess_arr, _pad_rows(agraph), _pad_rows(bgraph), scope


def tensorize(batch: Sequence[SynthMol]):
    """(tree_tensors, graph_tensors) in the layout of ``MolGraph.tensorize`` (ggpm/mol_graph.py:199-236):
    tree_tensors = (fnode [Nt+1,2], fmess [Et+1,4], agraph, bgraph, cgraph, scope),
    graph_tensors = (fnode [Na+1], fmess [Ea+1,4], agraph, bgraph, scope); int32 arrays."""
    tree = tensorize_graph([m.motif_labels for m in batch], [m.tree_edges for m in batch])
    graph = tensorize_graph([m.atom_labels for m in batch], [m.bonds for m in batch])
    tree_scope, graph_scope = tree[-1], graph[-1]
    max_cls = max(len(c) for m in batch for c in m.clusters)
    cgraph = np.zeros((tree[0].shape[0], max_cls), dtype=np.int32)
    for b, m in enumerate(batch):
        t0, a0 = tree_scope[b][0], graph_scope[b][0]
        for i, cls in enumerate(m.clusters):
            cgraph[t0 + i, :len(cls)] = [a + a0 for a in cls]
    return tree[:4] + (cgraph, tree_scope), graph


def batch_stats(tree_tensors, graph_tensors):
    """Sizes used for the roofline arithmetic: rows and true neighbour counts (non-zero entries)."""
    tf, tm, ta, tb, tc, _ = tree_tensors
    gf, gm, ga, gb, _ = graph_tensors
    return dict(
        Na=int(gf.shape[0] - 1), Ea=int(gm.shape[0] - 1), Nt=int(tf.shape[0] - 1), Et=int(tm.shape[0] - 1),
        nnz_bA=int(np.count_nonzero(gb)), nnz_aA=int(np.count_nonzero(ga)),
        nnz_bT=int(np.count_nonzero(tb)), nnz_aT=int(np.count_nonzero(ta)), nnz_c=int(np.count_nonzero(tc)),
    )


class SizeOnlyPairVocab:
    """Motif vocabulary stand-in: the encoder constructor only calls .size() (ggpm/encoder.py:52,57);
    defaults are PairVocab(trio_polymer_vocab.txt).size() == (721, 6215)."""

    def __init__(self, n_motif=721, n_attach=6215):
        self._size = (n_motif, n_attach)

    def size(self):
        return self._size


class SizeOnlyAtomVocab:
    """common_atom_vocab stand-in (ggpm/vocab.py:64-69): 38 atom types."""

    def __init__(self, n=38):
        self._n = n

    def size(self):
        return self._n
