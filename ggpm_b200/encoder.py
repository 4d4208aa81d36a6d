"""Hierarchical message-passing encoder — drop-in for ggpm/encoder.py MPNEncoder (:8-38),
HierMPNEncoder (:41-157), MotifEncoder (:252-341) and the incremental encoders the decoder drives: IncMPNEncoder
(:160-179), IncHierMPNEncoder (:182-249), IncEncoder (:344-394).

Same constructor arguments (the configs/*.json keys rnn_type, embed_size, hidden_size, depthT, depthG,
dropout), same forward inputs/outputs, same parameter names / shapes / registration order (state_dicts
load strictly both ways), same tie_embedding contract.  All arithmetic on the path runs in
libggpm_b200.so; hidden-state tables are carried internally at a padded width Hp (multiple of 32) and
sliced to H on return.  nn.Dropout / nn.Embedding stay torch modules (they are tied to the decoder's
modules by tie_embedding and must keep the reference's RNG behaviour).
"""
from __future__ import annotations

import weakref

import torch
import torch.nn as nn

from . import functional as Fn
from ._lib import ACT_RELU, ACT_TANH
from .nnutils import ints_to_device, to_cuda
from .rnn import GRU, LSTM

N_BOND_TYPES = 4     # len(MolGraph.BOND_LIST), ggpm/mol_graph.py:14
MAX_POS = 20         # MolGraph.MAX_POS, ggpm/mol_graph.py:15


def _i32(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.int32).contiguous()


def _message_graph(fmess, bgraph, n_nodes, n_table_rows, k2_offset=None, guard=None) -> Fn.MessageGraph:
    """bgraph CSR + per-message source / one-hot column ids; prepared (and range-checked) on the host when the batch
    came through ingest.PackedBatch, else derived and checked here (fmess columns: src, dst, label, position —
    ggpm/mol_graph.py:262-270).  The ids index the node table (src) and the one-hot weight-column table (k1, k2)
    inside the cell kernels; the reference's index_select raises on the same violations (ggpm/encoder.py:103-106,
    114-116, 121-125)."""
    b = Fn.csr_from_padded(bgraph, fmess.shape[0], guard=guard)
    cols = getattr(fmess, "_ggpm_cols", None)
    if cols is not None:
        src, k1, k2 = cols
        if k2_offset is None:
            k2 = None
    else:
        if guard is not None:
            k1_hi = n_table_rows if k2_offset is None else k2_offset
            guard.check(fmess[:, 0], 0, n_nodes, "message source node (fmess[:, 0])")
            guard.check(fmess[:, 2], 0, k1_hi, "message label (fmess[:, 2])")
            if k2_offset is not None:
                guard.check(fmess[:, 3], 0, n_table_rows - k2_offset, "message position (fmess[:, 3])")
        src, k1 = _i32(fmess[:, 0]), _i32(fmess[:, 2])
        k2 = None if k2_offset is None else _i32(fmess[:, 3] + k2_offset)
    mg = Fn.MessageGraph(b, src, k1, k2, n_nodes, n_table_rows)
    mg._by_src = getattr(fmess, "_ggpm_by_src", None)
    return mg


def _embed(seq, ids):
    """E_c / E_i lookup (ggpm/encoder.py:98,111).  The reference calls the nn.Sequential(Embedding, Dropout) module —
    possibly the decoder's, after tie_embedding; a plain Embedding + Dropout pair is evaluated as the same lookup and the
    same dropout module, with the library's embedding backward; anything else is called as a module."""
    if (isinstance(seq, nn.Sequential) and len(seq) == 2 and type(seq[0]) is nn.Embedding and type(seq[1]) is nn.Dropout
            and seq[0].padding_idx is None and seq[0].max_norm is None and not seq[0].scale_grad_by_freq
            and not seq[0].sparse and seq[0].weight.dtype == torch.float32 and ids.dim() == 1):
        return seq[1](Fn.embedding_rows(seq[0].weight, ids))
    return seq(ids)


def _public(t: torch.Tensor, H: int) -> torch.Tensor:
    """padded internal table -> fresh contiguous [rows, H] tensor"""
    return t if t.shape[1] == H else t[:, :H].contiguous()


class MPNEncoder(nn.Module):

    def __init__(self, rnn_type, input_size, node_fdim, hidden_size, depth, dropout):
        super(MPNEncoder, self).__init__()
        self.hidden_size = hidden_size
        self.input_size = input_size
        self.depth = depth
        self.W_o = nn.Sequential(
            nn.Linear(node_fdim + hidden_size, hidden_size),
            nn.ReLU(),
            nn.Dropout(dropout)
        )
        if rnn_type == 'GRU':
            self.rnn = GRU(input_size, hidden_size, depth)
        elif rnn_type == 'LSTM':
            self.rnn = LSTM(input_size, hidden_size, depth)
        else:
            raise ValueError('unsupported rnn cell type ' + rnn_type)

    # -- factored path used by the hierarchical encoders ------------------------------------------------
    def encode(self, phi_n, phi_x, mg: Fn.MessageGraph, a_csr: Fn.CSR, n_feat_x: int, n_bias_rows: int):
        """phi_n [N1, >=F]: node features for the read-out; phi_x [N1, >=Fx]: node features feeding the
        message inputs x_e = [phi_x[src_e] | one-hots].  Returns padded (node [N1, Hp], h [E1, Hp],
        incoming [N1, Hp]) where incoming[v] = sum of final messages into v."""
        H = self.hidden_size
        F = self.W_o[0].in_features - H
        out = self.rnn.messages(phi_x, mg, n_feat_x, n_bias_rows)
        h = out[0] if isinstance(out, tuple) else out
        incoming = Fn.gather_sum(h, a_csr)                                    # ggpm/encoder.py:31-32
        w = self.W_o[0].weight
        node = Fn.linear_act_split([phi_n, incoming], w, [F, H], self.W_o[0].bias,
                                   act=ACT_RELU, zero_row0=True)              # :33-38 (mask commutes with dropout)
        node = self.W_o[2](node)
        return node, h, incoming

    # -- reference signature (dense message inputs) -----------------------------------------------------
    def forward(self, fnode, fmess, agraph, bgraph):
        from .rnn import _identity_graph
        if fnode.shape[0] != agraph.shape[0]:
            raise ValueError("fnode and agraph disagree on the number of nodes")
        guard = Fn.IndexGuard(fmess.device)
        mg = _identity_graph(fmess, bgraph, guard)
        a_csr = Fn.csr_from_padded(agraph, fmess.shape[0], guard=guard)
        guard.raise_if_bad()
        H = self.hidden_size
        F = self.W_o[0].in_features - H
        out = self.rnn.messages(fmess, mg, self.input_size, 1)
        h = out[0] if isinstance(out, tuple) else out
        incoming = Fn.gather_sum(h, a_csr)
        w = self.W_o[0].weight
        node = Fn.linear_act_split([fnode, incoming], w, [F, H], self.W_o[0].bias, act=ACT_RELU, zero_row0=True)
        return self.W_o[2](_public(node, H)), _public(h, H)


class _TreeBatch:
    """Per-batch index structures shared by the attachment and motif layers."""

    def __init__(self, tree_tensors, guard=None):
        fnode, fmess, agraph, bgraph, cgraph, scope = tree_tensors
        self.n_nodes, self.n_mess = fnode.shape[0], fmess.shape[0]
        self.fnode = fnode
        self.cgraph = cgraph
        self.a = Fn.csr_from_padded(agraph, self.n_mess, guard=guard)
        self.mg = _message_graph(fmess, bgraph, self.n_nodes, MAX_POS, guard=guard)
        roots = getattr(scope, "roots_dev", None)
        if roots is None or roots.device != fnode.device:
            starts = [st for st, le in scope]
            if starts and not (0 <= min(starts) and max(starts) < self.n_nodes):
                raise IndexError("malformed batch: scope start outside [0, %d)" % self.n_nodes)
            roots = ints_to_device(starts, fnode.device)
        self.roots = roots


def _root_vectors(W_root, feat, incoming, roots, H):
    """W_root([feat[root] | sum of final messages into root]) with tanh, ggpm/encoder.py:128-138"""
    w = W_root[0].weight
    f = feat.index_select(0, roots)
    m = incoming.index_select(0, roots)
    return _public(Fn.linear_act_split([f, m], w, [H, H], W_root[0].bias, act=ACT_TANH), H)


class HierMPNEncoder(nn.Module):

    def __init__(self, vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout):
        super(HierMPNEncoder, self).__init__()
        self.vocab = vocab
        self.hidden_size = hidden_size
        self.dropout = dropout
        self.atom_size = atom_size = avocab.size()
        self.bond_size = bond_size = N_BOND_TYPES + MAX_POS

        self.E_c = nn.Sequential(nn.Embedding(vocab.size()[0], embed_size), nn.Dropout(dropout))
        self.E_i = nn.Sequential(nn.Embedding(vocab.size()[1], embed_size), nn.Dropout(dropout))
        self.W_c = nn.Sequential(nn.Linear(embed_size + hidden_size, hidden_size), nn.ReLU(), nn.Dropout(dropout))
        self.W_i = nn.Sequential(nn.Linear(embed_size * 2, hidden_size), nn.ReLU(), nn.Dropout(dropout))

        # plain tensors, not buffers: absent from state_dict and re-bindable by tie_embedding (ggpm/encoder.py:74-77)
        self.E_a = to_cuda(torch.eye(atom_size))
        self.E_b = to_cuda(torch.eye(N_BOND_TYPES))
        self.E_apos = to_cuda(torch.eye(MAX_POS))
        self.E_pos = to_cuda(torch.eye(MAX_POS))

        self.W_root = nn.Sequential(nn.Linear(hidden_size * 2, hidden_size), nn.Tanh())
        self.tree_encoder = MPNEncoder(rnn_type, hidden_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)
        self.inter_encoder = MPNEncoder(rnn_type, hidden_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)
        self.graph_encoder = MPNEncoder(rnn_type, atom_size + bond_size, atom_size, hidden_size, depthG, dropout)

    def tie_embedding(self, other):
        self.E_c, self.E_i = other.E_c, other.E_i
        self.E_a, self.E_b = other.E_a, other.E_b

    def forward(self, tree_tensors, graph_tensors):
        H = self.hidden_size
        g_fnode, g_fmess, g_agraph, g_bgraph, _ = graph_tensors
        guard = Fn.IndexGuard(g_fnode.device)
        tb = _TreeBatch(tree_tensors, guard)

        # atom layer: x_e = [E_a[atom(src_e)] | E_b[bond_e] | E_apos[pos_e]]  (embed_graph, :119-126)
        n_atoms, n_bonds = g_fnode.shape[0], g_fmess.shape[0]
        pad = (-self.atom_size) % 4                       # 16-byte rows: both uses below read it at this width
        atom_feat = nn.functional.pad(self.E_a, (0, pad)).index_select(0, g_fnode)
        mg = _message_graph(g_fmess, g_bgraph, n_atoms, N_BOND_TYPES + MAX_POS, k2_offset=N_BOND_TYPES, guard=guard)
        a_csr = Fn.csr_from_padded(g_agraph, n_bonds, guard=guard)
        c_csr = Fn.csr_from_padded(tb.cgraph, n_atoms, guard=guard)
        guard.raise_if_bad()                              # one read-back per forward, before any consumer of the ids runs
        hatom, _, _ = self.graph_encoder.encode(atom_feat, atom_feat, mg, a_csr, self.atom_size, N_BOND_TYPES)

        # attachment layer (embed_inter, :96-107)
        attach_emb = _embed(self.E_i, tb.fnode[:, 1])
        atoms_of_motif = Fn.gather_sum(hatom, c_csr)
        wi = self.W_i[0].weight
        E = attach_emb.shape[1]
        inter_feat = Fn.linear_act_split([attach_emb, atoms_of_motif], wi, [E, wi.shape[1] - E], self.W_i[0].bias,
                                         act=ACT_RELU)
        inter_feat = self.W_i[2](inter_feat)
        hinter, _, _ = self.inter_encoder.encode(inter_feat, inter_feat, tb.mg, tb.a, H, MAX_POS)

        # motif layer (embed_tree, :109-117)
        motif_emb = _embed(self.E_c, tb.fnode[:, 0])
        wc = self.W_c[0].weight
        E = motif_emb.shape[1]
        tree_feat = Fn.linear_act_split([motif_emb, hinter], wc, [E, wc.shape[1] - E], self.W_c[0].bias,
                                        act=ACT_RELU)
        tree_feat = self.W_c[2](tree_feat)
        hnode, _, incoming = self.tree_encoder.encode(tree_feat, tree_feat, tb.mg, tb.a, H, MAX_POS)

        hroot = _root_vectors(self.W_root, tree_feat, incoming, tb.roots, H)
        return hroot, _public(hnode, H), _public(hinter, H), _public(hatom, H)


class MotifEncoder(nn.Module):

    def __init__(self, vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout):
        super(MotifEncoder, self).__init__()
        self.vocab = vocab
        self.hidden_size = hidden_size
        self.dropout = dropout
        self.atom_size = atom_size = avocab.size()
        self.bond_size = bond_size = N_BOND_TYPES + MAX_POS

        self.E_c = nn.Sequential(nn.Embedding(vocab.size()[0], embed_size), nn.Dropout(dropout))
        self.E_i = nn.Sequential(nn.Embedding(vocab.size()[1], embed_size), nn.Dropout(dropout))
        self.W_root = nn.Sequential(nn.Linear(hidden_size * 2, hidden_size), nn.Tanh())
        self.E_a = to_cuda(torch.eye(atom_size))
        self.E_pos = to_cuda(torch.eye(MAX_POS))
        self.tree_encoder = MPNEncoder(rnn_type, hidden_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)

    def tie_embedding(self, other):
        self.E_c, self.E_i = other.E_c, other.E_i
        self.E_a = other.E_a

    def forward(self, tree_tensors):
        H = self.hidden_size
        guard = Fn.IndexGuard(tree_tensors[0].device)
        tb = _TreeBatch(tree_tensors, guard)
        guard.raise_if_bad()
        motif_emb = _embed(self.E_c, tb.fnode[:, 0])         # node feature of the read-out (embed_tree, :305-317)
        attach_emb = _embed(self.E_i, tb.fnode[:, 1])        # feeds the message inputs
        node, _, incoming = self.tree_encoder.encode(motif_emb, attach_emb, tb.mg, tb.a, H, MAX_POS)
        root = _root_vectors(self.W_root, motif_emb, incoming, tb.roots, H)
        return root, _public(node, H)


# ----------------------------------------------------------------------------------------------------------------
# incremental encoders driven by the decoder: ggpm/encoder.py:160-249, 344-394
# ----------------------------------------------------------------------------------------------------------------
_PADDED_LINEARS = weakref.WeakKeyDictionary()     # nn.Linear -> (key, (weight, bias) zero padded to 32-row multiples)


def _padded_linear(lin):
    """(W [Np, K], b [Np]) of an nn.Linear with the output rows zero padded to Np = multiple of 32, shared between the
    decoder steps while the parameters are unchanged (same key as rnn._GraphRNN._packed_for_steps).  With N = 250 the
    unpadded layer keeps its two gradient GEMMs on the CUDA-core engine (N % 4 != 0); padded, all three run on tcgen05
    and the ≈ 37 uses per batch share one padding node, so the parameter receives one summed gradient."""
    import torch.nn.functional as F_
    w, b = lin.weight, lin.bias
    key = (torch.is_grad_enabled(), w.data_ptr(), w._version, w.requires_grad, w.device,
           None if b is None else (b.data_ptr(), b._version, b.requires_grad))
    cached = _PADDED_LINEARS.get(lin)
    if cached is None or cached[0] != key:
        pad = Fn.round_up(w.shape[0], 32) - w.shape[0]
        cached = (key, (F_.pad(w, (0, 0, 0, pad)), None if b is None else F_.pad(b, (0, pad))))
        _PADDED_LINEARS[lin] = cached
    return cached[1]


def _linear_seq(seq, parts, widths):
    """nn.Sequential(Linear, ReLU, Dropout) on the concatenation of `parts`.  The incremental steps touch few rows, so
    the reference's own shape is the cheapest: one small torch.cat and ONE GEMM against the unsplit weight (its row
    length F + H is a multiple of 4 in every layer of the path, which keeps it on the tensor-core engine); the output
    rows are padded to a multiple of 32 columns (zero weights) and returned as the [:, :H] view."""
    lin = seq[0]
    H = lin.out_features
    if sum(widths) % 4 == 0:
        A = torch.cat([p[:, :k] for p, k in zip(parts, widths)], dim=1)
        W, b = _padded_linear(lin)
        out = Fn.linear_act([A], [W], b, act=ACT_RELU)
    else:
        out = Fn.linear_act_split(parts, lin.weight, widths, lin.bias, act=ACT_RELU)
    return seq[2](out if out.shape[1] == H else out[:, :H])


class IncMPNEncoder(MPNEncoder):
    """ggpm/encoder.py:160-179: recompute the sub-messages `submess` of the state table h, then the read-out of the
    nodes `subnode`; returns (node table [num_nodes, H] that is zero outside subnode, new state table(s))."""

    def __init__(self, rnn_type, input_size, node_fdim, hidden_size, depth, dropout):
        super(IncMPNEncoder, self).__init__(rnn_type, input_size, node_fdim, hidden_size, depth, dropout)

    def forward(self, tensors, h, num_nodes, subset):
        fnode, fmess, agraph, bgraph = tensors
        subnode, submess = subset
        if len(submess) > 0:
            h = self.rnn.sparse_forward(h, fmess, submess, bgraph)
        nei_message = Fn.slot_gather_sum(self.rnn.get_hidden_state(h), agraph)
        F = self.W_o[0].in_features - self.hidden_size
        node_hiddens = _linear_seq(self.W_o, [fnode, nei_message], [F, self.hidden_size])
        return Fn.scatter_into_zeros(node_hiddens, subnode, num_nodes), h      # index_scatter into a zero node_buf


def _check_tables_once(guard, tensors, n_mess_rows, n_atom_rows=None):
    """range-check the neighbour tables of one batch the first time an incremental encoder sees them (they are the
    same tensor objects for every decoder step of the batch; the row subsets taken from them per step go through
    torch's own index_select, which asserts its indices): agraph / bgraph entries index the message state table
    (n_mess_rows = its row count, appended per-molecule rows included), cgraph entries the atom table"""
    fnode, fmess, agraph, bgraph = tensors[:4]
    if getattr(bgraph, "_ggpm_checked", None) == (n_mess_rows, n_atom_rows):
        return
    for t, hi, what in ((agraph, n_mess_rows, "agraph"), (bgraph, n_mess_rows, "bgraph")):
        if isinstance(t, torch.Tensor) and t.dtype in (torch.int32, torch.int64) and t.numel():
            guard.check(t.reshape(-1), 0, hi, what)
    if len(tensors) == 6 and n_atom_rows is not None and isinstance(tensors[4], torch.Tensor) and tensors[4].numel():
        guard.check(tensors[4].reshape(-1), 0, n_atom_rows, "cgraph")
    try:
        bgraph._ggpm_checked = (n_mess_rows, n_atom_rows)
    except Exception:
        pass


def _sub_tensors(tensors, subset):
    """rows of (fnode, fmess, agraph, bgraph[, cgraph]) selected by (subnode, submess) — ggpm/encoder.py:197-208"""
    subnode, submess = subset
    fnode, fmess, agraph, bgraph = tensors[:4]
    fnode, fmess = fnode.index_select(0, subnode), fmess.index_select(0, submess)
    agraph, bgraph = agraph.index_select(0, subnode), bgraph.index_select(0, submess)
    if len(tensors) == 6:
        return fnode, fmess, agraph, bgraph, tensors[4].index_select(0, subnode), tensors[-1]
    return fnode, fmess, agraph, bgraph, tensors[-1]


def _sub_message_inputs(hnode, fmess, subnode, num_nodes, E_pos):
    """x of the sub-messages: [features of the source node | one-hot position] (ggpm/encoder.py:224-229); only nodes in
    `subnode` carry features — any other source reads a zero row, as the reference's scattered node_buf does"""
    node_buf = Fn.scatter_into_zeros(hnode, subnode, num_nodes)
    hmess = Fn.slot_gather_sum(node_buf, fmess[:, 0])
    return torch.cat([hmess, E_pos.index_select(0, fmess[:, 2])], dim=-1)


class IncHierMPNEncoder(HierMPNEncoder):

    def __init__(self, vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout):
        super(IncHierMPNEncoder, self).__init__(vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout)
        self.tree_encoder = IncMPNEncoder(rnn_type, hidden_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)
        self.inter_encoder = IncMPNEncoder(rnn_type, hidden_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)
        self.graph_encoder = IncMPNEncoder(rnn_type, self.atom_size + self.bond_size, self.atom_size, hidden_size, depthG,
                                           dropout)
        del self.W_root

    def embed_graph(self, graph_tensors):
        """ggpm/encoder.py:119-126 (one-hot rows; the decoder embeds the atom graph once per batch, decoder.py:193)"""
        fnode, fmess, agraph, bgraph, _ = graph_tensors
        hnode = self.E_a.index_select(0, fnode)
        hmess = torch.cat([hnode.index_select(0, fmess[:, 0]), self.E_b.index_select(0, fmess[:, 2]),
                           self.E_apos.index_select(0, fmess[:, 3])], dim=-1)
        return hnode, hmess, agraph, bgraph

    def get_sub_tensor(self, tensors, subset):
        return _sub_tensors(tensors, subset)

    def embed_sub_tree(self, tree_tensors, hinput, subtree, is_inter_layer):
        subnode, submess = subtree
        num_nodes = tree_tensors[0].size(0)
        fnode, fmess, agraph, bgraph, cgraph, _ = _sub_tensors(tree_tensors, subtree)
        if is_inter_layer:
            finput = _embed(self.E_i, fnode[:, 1])
            E = finput.shape[1]
            hinput = Fn.slot_gather_sum(hinput, cgraph)
            hnode = _linear_seq(self.W_i, [finput, hinput], [E, self.W_i[0].in_features - E])
        else:
            finput = _embed(self.E_c, fnode[:, 0])
            E = finput.shape[1]
            hinput = Fn.slot_gather_sum(hinput, subnode)
            hnode = _linear_seq(self.W_c, [finput, hinput], [E, self.W_c[0].in_features - E])
        hmess = fmess if len(submess) == 0 else _sub_message_inputs(hnode, fmess, subnode, num_nodes, self.E_pos)
        return hnode, hmess, agraph, bgraph

    def forward(self, tree_tensors, inter_tensors, graph_tensors, htree, hinter, hgraph, subtree, subgraph):
        num_tree_nodes = tree_tensors[0].size(0)
        num_graph_nodes = graph_tensors[0].size(0)
        guard = Fn.IndexGuard(tree_tensors[0].device)
        rows = lambda state: self.graph_encoder.rnn.get_hidden_state(state).shape[0]
        _check_tables_once(guard, graph_tensors, rows(hgraph.mess))
        _check_tables_once(guard, inter_tensors, rows(hinter.mess), num_graph_nodes)
        _check_tables_once(guard, tree_tensors, rows(htree.mess))
        guard.raise_if_bad()                               # a read-back only on the first step of a batch
        if len(subgraph[0]) + len(subgraph[1]) > 0:
            sub_graph_tensors = _sub_tensors(graph_tensors, subgraph)[:-1]     # already embedded by embed_graph
            hgraph.node, hgraph.mess = self.graph_encoder(sub_graph_tensors, hgraph.mess, num_graph_nodes, subgraph)
        if len(subtree[0]) + len(subtree[1]) > 0:
            sub_inter_tensors = self.embed_sub_tree(inter_tensors, hgraph.node, subtree, is_inter_layer=True)
            hinter.node, hinter.mess = self.inter_encoder(sub_inter_tensors, hinter.mess, num_tree_nodes, subtree)
            sub_tree_tensors = self.embed_sub_tree(tree_tensors, hinter.node, subtree, is_inter_layer=False)
            htree.node, htree.mess = self.tree_encoder(sub_tree_tensors, htree.mess, num_tree_nodes, subtree)
        return htree, hinter, hgraph


class IncEncoder(MotifEncoder):

    def __init__(self, vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout):
        super(IncEncoder, self).__init__(vocab, avocab, rnn_type, embed_size, hidden_size, depthT, depthG, dropout)
        self.tree_encoder = IncMPNEncoder(rnn_type, embed_size + MAX_POS, hidden_size, hidden_size, depthT, dropout)
        del self.W_root

    def get_sub_tensor(self, tensors, subset):
        return _sub_tensors(tensors, subset)

    def embed_sub_tree(self, tree_tensors, subtree):
        """ggpm/encoder.py:366-386: node features AND message inputs are the motif embedding E_c here"""
        subnode, submess = subtree
        num_nodes = tree_tensors[0].size(0)
        fnode, fmess, agraph, bgraph, cgraph, _ = _sub_tensors(tree_tensors, subtree)
        hnode = _embed(self.E_c, fnode[:, 0])
        hmess = fmess if len(submess) == 0 else _sub_message_inputs(hnode, fmess, subnode, num_nodes, self.E_pos)
        return hnode, hmess, agraph, bgraph

    def forward(self, tree_tensors, htree, subtree):
        num_tree_nodes = tree_tensors[0].size(0)
        guard = Fn.IndexGuard(tree_tensors[0].device)
        _check_tables_once(guard, tree_tensors, self.tree_encoder.rnn.get_hidden_state(htree.mess).shape[0])
        guard.raise_if_bad()
        if len(subtree[0]) + len(subtree[1]) > 0:
            sub_tree_tensors = self.embed_sub_tree(tree_tensors, subtree)
            htree.node, htree.mess = self.tree_encoder(sub_tree_tensors, htree.mess, num_tree_nodes, subtree)
        return htree
