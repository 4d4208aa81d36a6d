"""TEST INFRASTRUCTURE ONLY — CPU oracle for the incremental encoders the decoder drives.

Plain-torch restatement (dtype follows the weights; fp64 weights give an fp64 oracle) of

  * ``index_scatter``                           ggpm/nnutils.py:124-128
  * ``GRU.sparse_forward``                      ggpm/rnn.py:52-59
  * ``LSTM.sparse_forward``                     ggpm/rnn.py:110-121
  * ``IncMPNEncoder.forward``                   ggpm/encoder.py:160-179
  * ``IncHierMPNEncoder`` (get_sub_tensor, embed_sub_tree, forward)   ggpm/encoder.py:182-249
  * ``IncEncoder`` (embed_sub_tree, forward)    ggpm/encoder.py:344-394

over a reference-keyed ``state_dict``, plus a driver (``run_decode``) that feeds an incremental encoder the way the
decoder's teacher-forced loop does (ggpm/decoder.py:72-120 masks and initial state, :166-262 step loop): per step a
set of motif nodes / tree messages and the atoms / bonds of the motif just attached, with edge masks zeroing slots in
the middle of index rows.  The driver is written against the reference's module interface (``enc(...)``,
``enc.embed_graph``, ``enc.tree_encoder.rnn.get_init_state``), so the SAME function drives the unmodified reference,
this oracle (through the adapter classes below) and the CUDA product in the tests.

PARITY PIN: the reference has no tests for this path; ``oracle/make_golden_inc.py`` runs the unmodified reference
classes from /root/reference through ``run_decode`` (fp64 and fp32), asserts this oracle agrees to <= 1e-12 relative
in fp64, and commits the inputs / outputs / gradients under tests/golden_inc/.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Sequence

import torch

from .encoder_oracle import MAX_POS, N_BOND_TYPES, _affine, gather_padded

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------------
# restatement
# ----------------------------------------------------------------------------------------------
def index_scatter(sub: Tensor, table: Tensor, index: Tensor) -> Tensor:
    """ggpm/nnutils.py:124-128: rows `index` of `table` replaced by `sub` (mask-and-add in the reference)"""
    keep = torch.ones(table.shape[0], dtype=table.dtype)
    keep[index] = 0
    buf = torch.zeros_like(table).index_put((index,), sub)
    return table * keep[:, None] + buf


def _zero_rows(table: Tensor, index: Tensor) -> Tensor:
    keep = torch.ones(table.shape[0], dtype=table.dtype)
    keep[index] = 0
    return table * keep[:, None]


def lstm_cell(sd, p: str, x: Tensor, h_nb: Tensor, c_nb: Tensor):
    """ggpm/rnn.py:85-94"""
    xs = torch.cat([x, h_nb.sum(1)], dim=-1)
    gi = torch.sigmoid(_affine(sd, p + "W_i.0", xs))
    go = torch.sigmoid(_affine(sd, p + "W_o.0", xs))
    cand = torch.tanh(_affine(sd, p + "W.0", xs))
    x_rep = x[:, None, :].expand(-1, h_nb.shape[1], -1)
    gf = torch.sigmoid(_affine(sd, p + "W_f.0", torch.cat([x_rep, h_nb], dim=-1)))
    c = gi * cand + (gf * c_nb).sum(1)
    return go * torch.tanh(c), c


def gru_cell(sd, p: str, x: Tensor, h_nb: Tensor):
    """ggpm/rnn.py:25-39"""
    s = h_nb.sum(1)
    z = torch.sigmoid(_affine(sd, p + "W_z", torch.cat([x, s], dim=-1)))
    r = torch.sigmoid(_affine(sd, p + "W_r", x)[:, None, :] + _affine(sd, p + "U_r", h_nb))
    sg = (r * h_nb).sum(1)
    cand = torch.tanh(_affine(sd, p + "W_h", torch.cat([x, sg], dim=-1)))
    return (1.0 - z) * s + z * cand


def sparse_forward(sd, p: str, rnn_type: str, state, x: Tensor, submess: Tensor, bgraph: Tensor, depth: int):
    """ggpm/rnn.py:52-59 (GRU: state = h) and :110-121 (LSTM: state = (h, c))"""
    if rnn_type == "LSTM":
        h, c = state
        h, c = _zero_rows(h, submess), _zero_rows(c, submess)
        for _ in range(depth):
            sub_h, sub_c = lstm_cell(sd, p, x, gather_padded(h, bgraph), gather_padded(c, bgraph))
            h, c = index_scatter(sub_h, h, submess), index_scatter(sub_c, c, submess)
        return h, c
    h = _zero_rows(state, submess)
    for _ in range(depth):
        h = index_scatter(gru_cell(sd, p, x, gather_padded(h, bgraph)), h, submess)
    return h


def inc_mpn(sd, p: str, rnn_type: str, depth: int, tensors, state, num_nodes: int, subset):
    """IncMPNEncoder.forward, ggpm/encoder.py:165-179"""
    fnode, fmess, agraph, bgraph = tensors
    subnode, submess = subset
    if len(submess) > 0:
        state = sparse_forward(sd, p + "rnn.", rnn_type, state, fmess, submess, bgraph, depth)
    hid = state[0] if rnn_type == "LSTM" else state
    nei = gather_padded(hid, agraph).sum(1)
    node = torch.relu(_affine(sd, p + "W_o.0", torch.cat([fnode, nei], dim=-1)))
    buf = torch.zeros(num_nodes, node.shape[1], dtype=node.dtype)
    return index_scatter(node, buf, subnode), state


def sub_tensors(tensors, subset):
    """get_sub_tensor, ggpm/encoder.py:197-208"""
    subnode, submess = subset
    fnode, fmess, agraph, bgraph = tensors[:4]
    out = (fnode.index_select(0, subnode), fmess.index_select(0, submess), agraph.index_select(0, subnode),
           bgraph.index_select(0, submess))
    if len(tensors) == 6:
        return out + (tensors[4].index_select(0, subnode), tensors[-1])
    return out + (tensors[-1],)


def _init_state(rnn_type: str, hidden: int, fmess: Tensor, init, dtype):
    """get_init_state, ggpm/rnn.py:18-20 / 74-80"""
    h = torch.zeros(len(fmess), hidden, dtype=dtype)
    if rnn_type != "LSTM":
        return h if init is None else torch.cat((h, init), dim=0)
    c = torch.zeros_like(h)
    if init is not None:
        h = torch.cat((h, init), dim=0)
        c = torch.cat((c, torch.zeros_like(init)), dim=0)
    return h, c


class _RnnAdapter:
    def __init__(self, rnn_type, hidden, dtype):
        self.rnn_type, self.hidden, self.dtype = rnn_type, hidden, dtype

    def get_init_state(self, fmess, init_state=None):
        return _init_state(self.rnn_type, self.hidden, fmess, init_state, self.dtype)

    def get_hidden_state(self, h):
        return h[0] if self.rnn_type == "LSTM" else h


class IncHierOracle:
    """IncHierMPNEncoder (ggpm/encoder.py:182-249) over a state_dict; callable like the module"""

    def __init__(self, sd: Dict[str, Tensor], rnn_type: str, depthT: int, depthG: int):
        self.sd, self.rnn_type, self.depthT, self.depthG = sd, rnn_type, depthT, depthG
        w = sd["tree_encoder.W_o.0.weight"]
        self.hidden, self.dtype = w.shape[0], w.dtype
        self.n_atom = sd["graph_encoder.W_o.0.weight"].shape[1] - self.hidden
        self.tree_encoder = SimpleNamespace(rnn=_RnnAdapter(rnn_type, self.hidden, self.dtype))

    def embed_graph(self, graph_tensors):
        """ggpm/encoder.py:119-126"""
        fnode, fmess, agraph, bgraph, _ = graph_tensors
        eye = lambda n: torch.eye(n, dtype=self.dtype)
        hnode = eye(self.n_atom).index_select(0, fnode)
        hmess = torch.cat([hnode.index_select(0, fmess[:, 0]), eye(N_BOND_TYPES).index_select(0, fmess[:, 2]),
                           eye(MAX_POS).index_select(0, fmess[:, 3])], dim=-1)
        return hnode, hmess, agraph, bgraph

    def _embed_sub_tree(self, tree_tensors, hinput, subtree, inter: bool):
        """ggpm/encoder.py:210-231"""
        sd = self.sd
        subnode, submess = subtree
        num_nodes = tree_tensors[0].shape[0]
        fnode, fmess, agraph, bgraph, cgraph, _ = sub_tensors(tree_tensors, subtree)
        if inter:
            finput = sd["E_i.0.weight"].index_select(0, fnode[:, 1])
            hin = gather_padded(hinput, cgraph).sum(1)
            hnode = torch.relu(_affine(sd, "W_i.0", torch.cat([finput, hin], dim=-1)))
        else:
            finput = sd["E_c.0.weight"].index_select(0, fnode[:, 0])
            hnode = torch.relu(_affine(sd, "W_c.0", torch.cat([finput, hinput.index_select(0, subnode)], dim=-1)))
        if len(submess) == 0:
            return hnode, fmess, agraph, bgraph
        buf = index_scatter(hnode, torch.zeros(num_nodes, self.hidden, dtype=self.dtype), subnode)
        pos = torch.eye(MAX_POS, dtype=self.dtype).index_select(0, fmess[:, 2])
        return hnode, torch.cat([buf.index_select(0, fmess[:, 0]), pos], dim=-1), agraph, bgraph

    def __call__(self, tree_tensors, inter_tensors, graph_tensors, htree, hinter, hgraph, subtree, subgraph):
        """ggpm/encoder.py:233-249"""
        sd, rt = self.sd, self.rnn_type
        nt, ng = tree_tensors[0].shape[0], graph_tensors[0].shape[0]
        if len(subgraph[0]) + len(subgraph[1]) > 0:
            sub = sub_tensors(graph_tensors, subgraph)[:-1]
            hgraph.node, hgraph.mess = inc_mpn(sd, "graph_encoder.", rt, self.depthG, sub, hgraph.mess, ng, subgraph)
        if len(subtree[0]) + len(subtree[1]) > 0:
            sub = self._embed_sub_tree(inter_tensors, hgraph.node, subtree, True)
            hinter.node, hinter.mess = inc_mpn(sd, "inter_encoder.", rt, self.depthT, sub, hinter.mess, nt, subtree)
            sub = self._embed_sub_tree(tree_tensors, hinter.node, subtree, False)
            htree.node, htree.mess = inc_mpn(sd, "tree_encoder.", rt, self.depthT, sub, htree.mess, nt, subtree)
        return htree, hinter, hgraph


class IncMotifOracle:
    """IncEncoder (ggpm/encoder.py:344-394) over a state_dict"""

    def __init__(self, sd: Dict[str, Tensor], rnn_type: str, depthT: int):
        self.sd, self.rnn_type, self.depthT = sd, rnn_type, depthT
        w = sd["tree_encoder.W_o.0.weight"]
        self.hidden, self.dtype = w.shape[0], w.dtype
        self.tree_encoder = SimpleNamespace(rnn=_RnnAdapter(rnn_type, self.hidden, self.dtype))

    def __call__(self, tree_tensors, htree, subtree):
        subnode, submess = subtree
        if len(subnode) + len(submess) == 0:
            return htree
        num_nodes = tree_tensors[0].shape[0]
        fnode, fmess, agraph, bgraph, _, _ = sub_tensors(tree_tensors, subtree)
        hnode = self.sd["E_c.0.weight"].index_select(0, fnode[:, 0])
        if len(submess) > 0:
            buf = index_scatter(hnode, torch.zeros(num_nodes, self.hidden, dtype=self.dtype), subnode)
            pos = torch.eye(MAX_POS, dtype=self.dtype).index_select(0, fmess[:, 2])
            fmess = torch.cat([buf.index_select(0, fmess[:, 0]), pos], dim=-1)
        htree.node, htree.mess = inc_mpn(self.sd, "tree_encoder.", self.rnn_type, self.depthT,
                                         (hnode, fmess, agraph, bgraph), htree.mess, num_nodes, subtree)
        return htree


from ggpm_b200.decode_loop import HTuple, decode_schedule, run_decode          # noqa: E402,F401  (index bookkeeping only)


def cotangents(outs: Sequence[Tensor], seed: int):
    """seeded N(0,1) cotangents, one per output (generated in fp64 on the CPU, cast to each output's dtype / device)"""
    gen = torch.Generator().manual_seed(seed)
    return [torch.randn(tuple(o.shape), generator=gen, dtype=torch.float64).to(dtype=o.dtype, device=o.device) for o in outs]


def scalar_loss(outs: Sequence[Tensor], cots: Sequence[Tensor]):
    return sum((o * c).sum() for o, c in zip(outs, cots))
