"""Host-side stand-in for the decoder's teacher-forced loop, restricted to what drives the incremental encoders
(SURVEY 8(f) rank 1): the per-step node / message / atom / bond subsets of a synthetic batch (``decode_schedule``:
label_tree's DFS order, ggpm/mol_graph.py:120-139) and the loop itself (``run_decode``: initial state and edge masks of
ggpm/decoder.py:72-120, step loop of :166-262 without the prediction heads).  Index bookkeeping only — no arithmetic
on features happens here; ``enc`` may be this package's IncHierMPNEncoder / IncEncoder, the unmodified reference's, or
the oracle's adapters (tests), since all three expose the reference's module interface.  Used by the tests, the golden
generator and the timing tool of the incremental path; the decoder proper (prediction heads, attachment enumeration,
rdkit) is out of scope.
"""
from __future__ import annotations

from typing import List

import numpy as np
import torch

Tensor = torch.Tensor


def decode_schedule(mols, tree_tensors, graph_tensors):
    """Per-step work of the decoder's loop for a batch of ggpm_b200.synth.SynthMol, in batch-global ids.

    Order per molecule = label_tree's DFS (ggpm/mol_graph.py:120-139): (x, y, 1) going down to child y, (y, x, 0)
    coming back, then (root, None, 0).  Step t takes the t-th entry of every molecule that still has one
    (ggpm/decoder.py:201-212): subtree = ([x], [message x->y]); the atoms of y's cluster become the next step's new
    atoms and the bonds among them its new bonds (ggpm/decoder.py:84-100, 228-230, 260)."""
    t_fmess, t_scope = np.asarray(tree_tensors[1]), tree_tensors[-1]
    g_fmess, g_scope = np.asarray(graph_tensors[1]), graph_tensors[-1]
    t_mess = {(int(u), int(v)): i for i, (u, v) in enumerate(t_fmess[:, :2]) if i > 0}
    orders, clusters = [], {}
    for b, m in enumerate(mols):
        t0, a0 = t_scope[b][0], g_scope[b][0]
        nbrs = [[] for _ in m.motif_labels]
        for u, v, _ in m.tree_edges:
            nbrs[u].append(v)
        for i, cls in enumerate(m.clusters):
            clusters[t0 + i] = [a + a0 for a in cls]
        order = []

        def dfs(x, fa):
            for y in sorted(n for n in nbrs[x] if n != fa):
                order.append((t0 + x, t0 + y))
                dfs(y, x)
                order.append((t0 + y, t0 + x))

        dfs(0, -1)
        order.append((t0, None))
        orders.append(order)

    def induced(atoms):
        s = set(atoms)
        return [i for i, (u, v) in enumerate(g_fmess[:, :2]) if i > 0 and int(u) in s and int(v) in s]

    first_atoms = [a for b in range(len(mols)) for a in clusters[t_scope[b][0]]]
    root_out = [[t_mess[(t_scope[b][0], t_scope[b][0] + v)] for u, v, _ in m.tree_edges if u == 0]
                for b, m in enumerate(mols)]
    steps = []
    for t in range(max(len(o) for o in orders)):
        nodes, mess, atoms = [], [], []
        for o in orders:
            if t >= len(o):
                continue
            x, y = o[t]
            nodes.append(x)
            if y is not None:
                mess.append(t_mess[(x, y)])
                atoms.extend(clusters[y])
        steps.append(dict(nodes=nodes, mess=mess, next_atoms=atoms, next_bonds=induced(atoms)))
    return dict(first_atoms=first_atoms, first_bonds=induced(first_atoms), root_out=root_out, steps=steps)


class HTuple:
    """ggpm/decoder.py:13-16"""

    def __init__(self, node=None, mess=None, vmask=None, emask=None):
        self.node, self.mess, self.vmask, self.emask = node, mess, vmask, emask


def _mask_index(index: Tensor, mask: Tensor) -> Tensor:
    """index * index_select_ND(mask, 0, index)   (ggpm/decoder.py:72-83)"""
    return index * mask.index_select(0, index.reshape(-1)).reshape(index.shape)


def run_decode(enc, hier: bool, tree_tensors, graph_tensors, sched, init_vecs, device="cpu", max_steps=None):
    """Drive `enc` through the schedule.  tree_tensors / graph_tensors: torch.long tensors (+ scope list) on `device`.
    Returns a list of float tensors in a fixed order: per step the node rows, message rows and atom rows the decoder
    reads (topology / cluster / attachment predictions, ggpm/decoder.py:226-259), then the final state tables."""
    steps = sched["steps"] if max_steps is None else sched["steps"][:max_steps]
    # Every index list of the schedule goes to the device in ONE copy before the loop (teacher forcing: the order is
    # known up front).  A torch.tensor(list, device=...) per step, as ggpm/decoder.py:213-214 does, is a pageable
    # host-to-device copy that makes the host wait for the stream four times per step.
    root_out = sched["root_out"]
    lists = [sched["first_atoms"], sched["first_bonds"], [r for r, _ in tree_tensors[-1]],
             [m for ms in root_out for m in ms], [i for i, ms in enumerate(root_out) for _ in ms]]
    for st in steps:
        lists += [st["nodes"], st["mess"], st["next_atoms"], st["next_bonds"]]
    flat = torch.tensor([v for lst in lists for v in lst], dtype=torch.long)
    if torch.device(device).type == "cuda":
        flat = flat.pin_memory().to(device, non_blocking=True)
    offs = np.cumsum([0] + [len(lst) for lst in lists])
    staged = [flat[offs[i]:offs[i + 1]] for i in range(len(lists))]
    first_atoms, first_bonds, roots, root_mess, root_mess_mol = staged[:5]
    step_lists = [staged[5 + 4 * i: 9 + 4 * i] for i in range(len(steps))]
    fnode, fmess, agraph, bgraph, cgraph, scope = tree_tensors
    B, num_mess = len(scope), fmess.shape[0]
    rnn = enc.tree_encoder.rnn
    inter_tensors = list(tree_tensors)

    # init_decoder_state (ggpm/decoder.py:102-120): one extra state row per molecule, wired into the root's rows
    agraph, bgraph = agraph.clone(), bgraph.clone()
    agraph[roots, -1] = num_mess + torch.arange(B, dtype=agraph.dtype, device=agraph.device)
    bgraph[root_mess, -1] = num_mess + root_mess_mol.to(bgraph.dtype)
    tree_tensors = [fnode, fmess, agraph, bgraph, cgraph, scope]
    htree = HTuple(mess=rnn.get_init_state(fmess, init_vecs),
                   emask=torch.cat([bgraph.new_zeros(num_mess), bgraph.new_ones(B)], dim=0))
    outs: List[Tensor] = []
    if hier:
        hinter = HTuple(mess=rnn.get_init_state(fmess), emask=bgraph.new_zeros(num_mess))
        hgraph = HTuple(mess=rnn.get_init_state(graph_tensors[1]), vmask=bgraph.new_zeros(graph_tensors[0].shape[0]),
                        emask=bgraph.new_zeros(graph_tensors[1].shape[0]))
        atoms, bonds = first_atoms, first_bonds
        hgraph.vmask.scatter_(0, atoms, 1)
        hgraph.emask.scatter_(0, bonds, 1)
        subgraph = (atoms, bonds)
        graph_tensors = tuple(enc.embed_graph(graph_tensors)) + (graph_tensors[-1],)

    for nodes, mess, next_atoms, next_bonds in step_lists:
        subtree = (nodes, mess)
        htree.emask.scatter_(0, subtree[1], 1)
        t_f, t_m, t_a, t_b, t_c, t_s = tree_tensors
        if hier:
            hinter.emask.scatter_(0, subtree[1], 1)
            cur_tree = (t_f, t_m, _mask_index(t_a, htree.emask), _mask_index(t_b, htree.emask),
                        _mask_index(t_c, hgraph.vmask), t_s)
            i_f, i_m, i_a, i_b, i_c, i_s = inter_tensors
            cur_inter = (i_f, i_m, _mask_index(i_a, hinter.emask), _mask_index(i_b, hinter.emask),
                         _mask_index(i_c, hgraph.vmask), i_s)
            g_f, g_m, g_a, g_b, g_s = graph_tensors
            cur_graph = (g_f, g_m, _mask_index(g_a, hgraph.emask), _mask_index(g_b, hgraph.emask), g_s)
            htree, hinter, hgraph = enc(cur_tree, cur_inter, cur_graph, htree, hinter, hgraph, subtree, subgraph)
            outs.append(hgraph.node.index_select(0, subgraph[0]))
        else:
            cur_tree = (t_f, t_m, _mask_index(t_a, htree.emask), _mask_index(t_b, htree.emask), t_c, t_s)
            htree = enc(cur_tree, htree, subtree)
        outs.append(htree.node.index_select(0, subtree[0]))
        if len(subtree[1]) > 0:
            outs.append(rnn.get_hidden_state(htree.mess).index_select(0, subtree[1]))
        if hier:
            atoms, bonds = next_atoms, next_bonds
            if len(atoms) > 0:
                hgraph.vmask.scatter_(0, atoms, 1)
            if len(bonds) > 0:
                hgraph.emask.scatter_(0, bonds, 1)
            subgraph = (atoms, bonds)

    def tables(s):
        return list(s) if isinstance(s, tuple) else [s]

    outs += tables(htree.mess)
    if hier:
        outs += tables(hinter.mess) + tables(hgraph.mess)
    return outs
