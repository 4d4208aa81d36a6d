"""TEST INFRASTRUCTURE ONLY — CPU oracle for the hierarchical message-passing encoder.

A functional restatement (plain torch CPU ops over a reference-keyed ``state_dict``; dtype follows
the weights, so fp64 weights give an fp64 oracle) of the algorithm of

  * ``index_select_ND``                      ggpm/nnutils.py:65-70
  * ``LSTM.LSTM`` / ``LSTM.forward``         ggpm/rnn.py:85-108
  * ``GRU.GRU`` / ``GRU.forward``            ggpm/rnn.py:25-50
  * ``MPNEncoder.forward``                   ggpm/encoder.py:28-38
  * ``HierMPNEncoder.embed_*`` / ``forward`` ggpm/encoder.py:96-157
  * ``MotifEncoder.embed_*`` / ``forward``   ggpm/encoder.py:305-341

in the reference's own formulation: padded neighbour tensors are gathered in full (padding slots
read row 0, which is kept at zero), the gate inputs are concatenations ``[x | h]`` and every gate
is one affine map on that concatenation.  This is deliberately the expensive padded form, so that
timing it is timing the reference's CPU algorithm (bench.py ``cpu_baseline`` / ``--impl reference``).

Dropout layers are identities by default (eval mode or p = 0), which is the only regime where parity with
the reference is defined bit-for-bit.  For train-mode parity the encoders take ``drop``: a dict
``site -> multiplier`` (mask / (1 - p), shaped like the site's activation) applied exactly where the reference's
``nn.Dropout`` modules sit (E_i / E_c: ggpm/encoder.py:51-55,98,111; W_i / W_c: :60-72; W_o of each MPNEncoder:
:15-19); tests install the same fixed multipliers in place of the product's dropout modules.

ReLU kinks: ``relu`` (dict ``site -> 0/1 mask``) makes the five ReLU sites (W_i, W_c, W_o of each MPNEncoder) evaluate
``pre * mask`` instead of ``relu(pre)``, i.e. the gradient of a PRESCRIBED linear branch; ``trace`` (dict) receives the
pre-activations.  Full-size parity tests use it to compare gradients on the activation pattern the implementation
under test actually took, and to prove that every unit whose pattern differs from the fp64 one sits on a kink
(|pre-activation| at rounding level) — see tests/test_gpu_parity_full.py.

PARITY PIN: the reference ships no tests or golden vectors for this path, so the pin
is the reference itself executed in the build container: ``oracle/make_golden.py`` runs the
unmodified reference modules from /root/reference (fp64 and fp32) on seeded inputs, asserts this
oracle agrees with them to <= 1e-12 relative in fp64, and commits inputs/outputs/gradients under
tests/golden/.  tests/test_oracle_golden.py re-checks the oracle against those files everywhere,
and against the live reference whenever /root/reference exists.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

from typing import Dict, Sequence, Tuple

import torch

Tensor = torch.Tensor
N_BOND_TYPES = 4      # len(MolGraph.BOND_LIST), ggpm/mol_graph.py:14
MAX_POS = 20          # MolGraph.MAX_POS, ggpm/mol_graph.py:15


def gather_padded(table: Tensor, index: Tensor) -> Tensor:
    """[R, H] rows picked by a padded [M, K] index -> [M, K, H]  (ggpm/nnutils.py:65-70)."""
    picked = torch.index_select(table, 0, index.reshape(-1))
    return picked.reshape(tuple(index.shape) + tuple(table.shape[1:]))


def _affine(sd: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    w = sd[name + ".weight"]
    y = x @ w.t()
    b = sd.get(name + ".bias")
    return y if b is None else y + b


# Full-size comparisons (tests/test_gpu_parity_full.py) evaluate the oracle in fp64 on the GPU, where the padded [E, K, H]
# intermediates of every depth round would not fit if autograd kept them all: with this flag each round is recomputed
# in the backward pass (torch.utils.checkpoint) — same arithmetic, same results (tests/test_oracle_golden.py).
CHECKPOINT_ROUNDS = False


def _round(fn, *state):
    if CHECKPOINT_ROUNDS and torch.is_grad_enabled():
        from torch.utils.checkpoint import checkpoint
        return checkpoint(fn, *state, use_reentrant=False)
    return fn(*state)


def _row0_mask(rows: int, like: Tensor) -> Tensor:
    m = torch.ones(rows, 1, dtype=like.dtype, device=like.device)
    m[0, 0] = 0
    return m


def lstm_messages(sd, prefix: str, x: Tensor, bgraph: Tensor, depth: int) -> Tuple[Tensor, Tensor]:
    """ggpm/rnn.py:96-108 with the cell of ggpm/rnn.py:85-94.  Keys: prefix + W_i.0 / W_o.0 / W_f.0 / W.0"""
    hidden = sd[prefix + "W_i.0.weight"].shape[0]
    h = torch.zeros(x.shape[0], hidden, dtype=x.dtype, device=x.device)
    c = torch.zeros_like(h)
    mask = _row0_mask(x.shape[0], x)

    def one_round(h, c):
        h_nb = gather_padded(h, bgraph)                   # [E, K, H]
        c_nb = gather_padded(c, bgraph)
        xs = torch.cat([x, h_nb.sum(1)], dim=-1)          # [E, IN + H]
        gate_i = torch.sigmoid(_affine(sd, prefix + "W_i.0", xs))
        gate_o = torch.sigmoid(_affine(sd, prefix + "W_o.0", xs))
        cand = torch.tanh(_affine(sd, prefix + "W.0", xs))
        x_rep = x[:, None, :].expand(-1, h_nb.shape[1], -1)
        gate_f = torch.sigmoid(_affine(sd, prefix + "W_f.0", torch.cat([x_rep, h_nb], dim=-1)))
        c = gate_i * cand + (gate_f * c_nb).sum(1)
        h = gate_o * torch.tanh(c)
        return h * mask, c * mask

    for _ in range(depth):
        h, c = _round(one_round, h, c)
    return h, c


def gru_messages(sd, prefix: str, x: Tensor, bgraph: Tensor, depth: int) -> Tensor:
    """ggpm/rnn.py:41-50 with the cell of ggpm/rnn.py:25-39.  Keys: prefix + W_z / W_r / U_r / W_h"""
    hidden = sd[prefix + "W_z.weight"].shape[0]
    h = torch.zeros(x.shape[0], hidden, dtype=x.dtype, device=x.device)
    mask = _row0_mask(x.shape[0], x)

    def one_round(h):
        h_nb = gather_padded(h, bgraph)
        s = h_nb.sum(1)
        z = torch.sigmoid(_affine(sd, prefix + "W_z", torch.cat([x, s], dim=-1)))
        r = torch.sigmoid(_affine(sd, prefix + "W_r", x)[:, None, :] + _affine(sd, prefix + "U_r", h_nb))
        sg = (r * h_nb).sum(1)
        cand = torch.tanh(_affine(sd, prefix + "W_h", torch.cat([x, sg], dim=-1)))
        h = (1.0 - z) * s + z * cand
        return h * mask

    for _ in range(depth):
        h = _round(one_round, h)
    return h


def _drop(drop, site: str, x: Tensor) -> Tensor:
    return x if not drop or site not in drop else x * drop[site].to(x.dtype)


def _relu(pre: Tensor, site: str, relu=None, trace=None) -> Tensor:
    if trace is not None:
        trace[site] = pre.detach()
    if relu is not None and site in relu:
        return pre * relu[site].to(pre.dtype)
    return torch.relu(pre)


def mpn_layer(sd, prefix: str, rnn_type: str, fnode: Tensor, fmess: Tensor, agraph: Tensor,
              bgraph: Tensor, depth: int, drop=None, relu=None, trace=None) -> Tuple[Tensor, Tensor]:
    """ggpm/encoder.py:28-38: message RNN, messages->nodes sum, W_o + ReLU, row-0 mask."""
    if rnn_type == "LSTM":
        h, _ = lstm_messages(sd, prefix + "rnn.", fmess, bgraph, depth)
    elif rnn_type == "GRU":
        h = gru_messages(sd, prefix + "rnn.", fmess, bgraph, depth)
    else:
        raise ValueError("unsupported rnn cell type " + rnn_type)
    incoming = gather_padded(h, agraph).sum(1)
    node = _relu(_affine(sd, prefix + "W_o.0", torch.cat([fnode, incoming], dim=-1)), prefix + "W_o", relu, trace)
    node = _drop(drop, prefix + "W_o", node)
    return node * _row0_mask(node.shape[0], node), h


def _onehot(index: Tensor, width: int, dtype) -> Tensor:
    return torch.eye(width, dtype=dtype, device=index.device).index_select(0, index)


def _root_vectors(sd, node_feat: Tensor, hmess: Tensor, agraph: Tensor, scope) -> Tensor:
    """ggpm/encoder.py:128-138 (and 319-328): W_root([feat[root] | sum of messages into root]), tanh."""
    roots = torch.tensor([st for st, _ in scope], dtype=torch.long, device=agraph.device)
    feat = node_feat.index_select(0, roots)
    incoming = gather_padded(hmess, agraph.index_select(0, roots)).sum(1)
    return torch.tanh(_affine(sd, "W_root.0", torch.cat([feat, incoming], dim=-1)))


def hier_encoder(sd: Dict[str, Tensor], rnn_type: str, depthT: int, depthG: int,
                 tree_tensors: Sequence, graph_tensors: Sequence, drop=None, relu=None, trace=None):
    """HierMPNEncoder.forward, ggpm/encoder.py:140-157.  Index tensors are torch.long (make_cuda,
    ggpm/nnutils.py:210-214).  Returns (hroot, hnode, hinter, hatom)."""
    dtype = sd["W_root.0.weight"].dtype
    t_fnode, t_fmess, t_agraph, t_bgraph, t_cgraph, t_scope = tree_tensors
    g_fnode, g_fmess, g_agraph, g_bgraph, _ = graph_tensors
    n_atom_types = sd["graph_encoder.W_o.0.weight"].shape[1] - sd["graph_encoder.W_o.0.weight"].shape[0]

    # atom layer (embed_graph, ggpm/encoder.py:119-126)
    atom_feat = _onehot(g_fnode, n_atom_types, dtype)
    bond_in = torch.cat([atom_feat.index_select(0, g_fmess[:, 0]),
                         _onehot(g_fmess[:, 2], N_BOND_TYPES, dtype),
                         _onehot(g_fmess[:, 3], MAX_POS, dtype)], dim=-1)
    hatom, _ = mpn_layer(sd, "graph_encoder.", rnn_type, atom_feat, bond_in, g_agraph, g_bgraph, depthG, drop, relu, trace)

    # attachment layer (embed_inter, ggpm/encoder.py:96-107)
    attach_emb = _drop(drop, "E_i", sd["E_i.0.weight"].index_select(0, t_fnode[:, 1]))
    atoms_of_motif = gather_padded(hatom, t_cgraph).sum(1)
    inter_feat = _relu(_affine(sd, "W_i.0", torch.cat([attach_emb, atoms_of_motif], dim=-1)), "W_i", relu, trace)
    inter_feat = _drop(drop, "W_i", inter_feat)
    pos = _onehot(t_fmess[:, 2], MAX_POS, dtype)
    inter_in = torch.cat([inter_feat.index_select(0, t_fmess[:, 0]), pos], dim=-1)
    hinter, _ = mpn_layer(sd, "inter_encoder.", rnn_type, inter_feat, inter_in, t_agraph, t_bgraph, depthT, drop, relu, trace)

    # motif layer (embed_tree, ggpm/encoder.py:109-117)
    motif_emb = _drop(drop, "E_c", sd["E_c.0.weight"].index_select(0, t_fnode[:, 0]))
    tree_feat = _relu(_affine(sd, "W_c.0", torch.cat([motif_emb, hinter], dim=-1)), "W_c", relu, trace)
    tree_feat = _drop(drop, "W_c", tree_feat)
    tree_in = torch.cat([tree_feat.index_select(0, t_fmess[:, 0]), pos], dim=-1)
    hnode, hmess = mpn_layer(sd, "tree_encoder.", rnn_type, tree_feat, tree_in, t_agraph, t_bgraph, depthT, drop, relu, trace)

    hroot = _root_vectors(sd, tree_feat, hmess, t_agraph, t_scope)
    return hroot, hnode, hinter, hatom


def motif_encoder(sd: Dict[str, Tensor], rnn_type: str, depthT: int, tree_tensors: Sequence, drop=None, relu=None,
                  trace=None):
    """MotifEncoder.forward, ggpm/encoder.py:330-341.  Returns (root, node)."""
    dtype = sd["W_root.0.weight"].dtype
    t_fnode, t_fmess, t_agraph, t_bgraph, _, t_scope = tree_tensors
    motif_emb = _drop(drop, "E_c", sd["E_c.0.weight"].index_select(0, t_fnode[:, 0]))
    attach_emb = _drop(drop, "E_i", sd["E_i.0.weight"].index_select(0, t_fnode[:, 1]))
    mess_in = torch.cat([attach_emb.index_select(0, t_fmess[:, 0]),
                         _onehot(t_fmess[:, 2], MAX_POS, dtype)], dim=-1)
    node, hmess = mpn_layer(sd, "tree_encoder.", rnn_type, motif_emb, mess_in, t_agraph, t_bgraph, depthT, drop, relu, trace)
    root = _root_vectors(sd, motif_emb, hmess, t_agraph, t_scope)
    return root, node


# ----------------------------------------------------------------------------------------------
# helpers shared by tests / bench (state_dict construction with the reference's shapes)
# ----------------------------------------------------------------------------------------------

def reference_state_shapes(kind: str, rnn_type: str, hidden: int, embed: int, n_motif=721,
                           n_attach=6215, n_atom=38):
    """Ordered {key: shape} of the reference module's state_dict (checked against the
    live reference in tests/test_oracle_golden.py).  kind in {"hier", "motif"}."""
    def rnn(prefix, inp):
        if rnn_type == "LSTM":
            out = []
            for g in ("W_i", "W_o", "W_f", "W"):
                out += [(f"{prefix}{g}.0.weight", (hidden, inp + hidden)), (f"{prefix}{g}.0.bias", (hidden,))]
            return out
        return [(prefix + "W_z.weight", (hidden, inp + hidden)), (prefix + "W_z.bias", (hidden,)),
                (prefix + "W_r.weight", (hidden, inp)),
                (prefix + "U_r.weight", (hidden, hidden)), (prefix + "U_r.bias", (hidden,)),
                (prefix + "W_h.weight", (hidden, inp + hidden)), (prefix + "W_h.bias", (hidden,))]

    def mpn(prefix, inp, fdim):
        return [(prefix + "W_o.0.weight", (hidden, fdim + hidden)), (prefix + "W_o.0.bias", (hidden,))] \
            + rnn(prefix + "rnn.", inp)

    keys = [("E_c.0.weight", (n_motif, embed)), ("E_i.0.weight", (n_attach, embed))]
    if kind == "hier":
        keys += [("W_c.0.weight", (hidden, embed + hidden)), ("W_c.0.bias", (hidden,)),
                 ("W_i.0.weight", (hidden, 2 * embed)), ("W_i.0.bias", (hidden,))]
    keys += [("W_root.0.weight", (hidden, 2 * hidden)), ("W_root.0.bias", (hidden,))]
    keys += mpn("tree_encoder.", hidden + MAX_POS, hidden)
    if kind == "hier":
        keys += mpn("inter_encoder.", hidden + MAX_POS, hidden)
        keys += mpn("graph_encoder.", n_atom + N_BOND_TYPES + MAX_POS, n_atom)
    return dict(keys)


def random_state(kind, rnn_type, hidden, embed, seed, dtype=torch.float32, bias_std=0.1, **vocab):
    """xavier_normal_ matrices as vae_train.py:49-53 does, but N(0, bias_std) biases (zero biases
    would hide bias bugs)."""
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    for key, shape in reference_state_shapes(kind, rnn_type, hidden, embed, **vocab).items():
        if len(shape) == 1:
            sd[key] = (torch.randn(shape, generator=gen, dtype=torch.float64) * bias_std).to(dtype)
        else:
            std = (2.0 / (shape[0] + shape[1])) ** 0.5
            sd[key] = (torch.randn(shape, generator=gen, dtype=torch.float64) * std).to(dtype)
    return sd
