"""Graph GRU / LSTM cells over a directed message graph — drop-in for ggpm/rnn.py (GRU :5-59,
LSTM :61-121), forward path.  Same constructor, attribute names, parameter names/shapes/registration
order (so state_dicts interchange), same forward(fmess, bgraph) contract; the arithmetic runs in
libggpm_b200.so (hoisted x-projection, CSR neighbour lists, hand-written BPTT)."""
from __future__ import annotations

import weakref

import torch
import torch.nn as nn

from . import functional as Fn
from . import packing


_STEP_PACKS = weakref.WeakKeyDictionary()      # module -> (key, packed operands), see _GraphRNN._packed_for_steps


def _identity_graph(fmess: torch.Tensor, bgraph: torch.Tensor, guard=None) -> Fn.MessageGraph:
    """dense message inputs: every message is its own 'node', one table row carries the bias.  Neighbour entries are
    range-checked (Fn.IndexGuard; read back here unless the caller passes its own guard and reads it back itself)."""
    E1 = fmess.shape[0]
    dev = fmess.device
    src = torch.arange(E1, dtype=torch.int32, device=dev)
    k1 = torch.zeros(E1, dtype=torch.int32, device=dev)
    own = guard is None
    if own:
        guard = Fn.IndexGuard(dev)
    csr = Fn.csr_from_padded(bgraph, E1, guard=guard)
    if own:
        guard.raise_if_bad()
    return Fn.MessageGraph(csr, src, k1, None, E1, 1)


class _GraphRNN(nn.Module):
    def __init__(self, input_size, hidden_size, depth):
        super().__init__()
        self.hidden_size = hidden_size
        self.input_size = input_size
        self.depth = depth

    @property
    def padded_hidden(self):
        return Fn.padded_width(self.hidden_size)

    def _packed_for_steps(self, pack):
        """Packed gate weights for sparse_forward, reused while the parameters are unchanged.  The decoder calls
        sparse_forward once per step (~37 times per batch) with the same weights; sharing ONE packing node between those
        calls means one pack launch per batch instead of one per step and, in the backward pass, one summed gradient
        per packed operand and one unpack.  The key holds every parameter's storage address and version counter
        (optimizer steps, load_state_dict and in-place initialisers bump it) plus the autograd mode."""
        params = list(self.parameters())
        key = (torch.is_grad_enabled(), self.padded_hidden) + tuple((p.data_ptr(), p._version, p.requires_grad, p.device)
                                                                    for p in params)
        cached = _STEP_PACKS.get(self)
        if cached is None or cached[0] != key:
            cached = (key, pack(self, self.input_size, self.padded_hidden, 1))
            _STEP_PACKS[self] = cached      # kept beside the module, not on it: deepcopy / pickle / state_dict never see it
        return cached[1]

    def _sparse_args(self, h, fmess, submess, bgraph):
        """validated (sub [M] int32, nbr [M*K] int32, K) of a sparse_forward call"""
        if fmess.dim() != 2 or fmess.shape[1] != self.input_size or fmess.dtype != torch.float32:
            raise ValueError("fmess must be float32 [M, %d], got %s %s" % (self.input_size, tuple(fmess.shape), fmess.dtype))
        if bgraph.dim() != 2 or bgraph.shape[0] != fmess.shape[0] or submess.numel() != fmess.shape[0]:
            raise ValueError("sparse_forward: fmess, submess and bgraph must describe the same sub-messages")
        if h.dim() != 2 or h.shape[1] != self.hidden_size or h.dtype != torch.float32:
            raise ValueError("state table must be float32 [R, %d]" % self.hidden_size)
        R = h.shape[0]
        return Fn.row_ids(submess.reshape(-1), R), Fn.row_ids(bgraph.reshape(-1), R), bgraph.shape[1]

    def _check(self, fmess, bgraph):
        if fmess.dim() != 2 or fmess.shape[1] != self.input_size:
            raise ValueError("fmess must be [E, %d], got %s" % (self.input_size, tuple(fmess.shape)))
        if bgraph.shape[0] != fmess.shape[0]:
            raise ValueError("bgraph and fmess disagree on the number of messages")


class GRU(_GraphRNN):

    def __init__(self, input_size, hidden_size, depth):
        super().__init__(input_size, hidden_size, depth)
        self.W_z = nn.Linear(input_size + hidden_size, hidden_size)
        self.W_r = nn.Linear(input_size, hidden_size, bias=False)
        self.U_r = nn.Linear(hidden_size, hidden_size)
        self.W_h = nn.Linear(input_size + hidden_size, hidden_size)

    def get_init_state(self, fmess, init_state=None):
        h = torch.zeros(len(fmess), self.hidden_size, device=fmess.device)
        return h if init_state is None else torch.cat((h, init_state), dim=0)

    def get_hidden_state(self, h):
        return h

    def messages(self, phi_x, mg: Fn.MessageGraph, n_feat: int, n_bias_rows: int):
        """factored entry used by the encoders: x_e = [phi_x[src_e] | one-hot(k1_e) | one-hot(k2_e)];
        returns the padded final state [E1, Hp]"""
        Hp = self.padded_hidden
        Wx, T, Uz, Ur, br, Uh = packing.pack_gru(self, n_feat, Hp, n_bias_rows)
        return Fn.GRUMessages.apply(packing.feature_columns(phi_x, n_feat), Wx, T, Uz, Ur, br, Uh, mg, self.depth,
                                    self.hidden_size)

    def forward(self, fmess, bgraph):
        self._check(fmess, bgraph)
        h = self.messages(fmess, _identity_graph(fmess, bgraph), self.input_size, 1)
        return h[:, :self.hidden_size].contiguous()

    def sparse_forward(self, h, fmess, submess, bgraph):
        """ggpm/rnn.py:52-59: rows `submess` of the table h [R, H] are reset and recomputed for `depth` rounds; fmess
        [M, IN] / bgraph [M, K] are the inputs / neighbour slots of those rows (slot values index the whole table)."""
        if submess.numel() == 0 or self.depth == 0:
            return h * 1.0 if submess.numel() == 0 else Fn.index_scatter(h.new_zeros(submess.numel(), h.shape[1]), h, submess)
        sub, nbr, K = self._sparse_args(h, fmess, submess, bgraph)
        Wx, T, Uz, Ur, br, Uh = self._packed_for_steps(packing.pack_gru)
        return Fn.SparseGRU.apply(h, packing.feature_columns(fmess, self.input_size), Wx, T, Uz, Ur, br, Uh, sub, nbr, K,
                                  self.depth)

class LSTM(_GraphRNN):

    def __init__(self, input_size, hidden_size, depth):
        super().__init__(input_size, hidden_size, depth)
        self.W_i = nn.Sequential(nn.Linear(input_size + hidden_size, hidden_size), nn.Sigmoid())
        self.W_o = nn.Sequential(nn.Linear(input_size + hidden_size, hidden_size), nn.Sigmoid())
        self.W_f = nn.Sequential(nn.Linear(input_size + hidden_size, hidden_size), nn.Sigmoid())
        self.W = nn.Sequential(nn.Linear(input_size + hidden_size, hidden_size), nn.Tanh())

    def get_init_state(self, fmess, init_state=None):
        h = torch.zeros(len(fmess), self.hidden_size, device=fmess.device)
        c = torch.zeros(len(fmess), self.hidden_size, device=fmess.device)
        if init_state is not None:
            h = torch.cat((h, init_state), dim=0)
            c = torch.cat((c, torch.zeros_like(init_state)), dim=0)
        return h, c

    def get_hidden_state(self, h):
        return h[0]

    def messages(self, phi_x, mg: Fn.MessageGraph, n_feat: int, n_bias_rows: int):
        Hp = self.padded_hidden
        Wx, T, U4 = packing.pack_lstm(self, n_feat, Hp, n_bias_rows)
        return Fn.LSTMMessages.apply(packing.feature_columns(phi_x, n_feat), Wx, T, U4, mg, self.depth,
                                     self.hidden_size)

    def forward(self, fmess, bgraph):
        self._check(fmess, bgraph)
        h, c = self.messages(fmess, _identity_graph(fmess, bgraph), self.input_size, 1)
        H = self.hidden_size
        return h[:, :H].contiguous(), c[:, :H].contiguous()

    def sparse_forward(self, h, fmess, submess, bgraph):
        """ggpm/rnn.py:110-121: as GRU.sparse_forward on the (h, c) pair of tables"""
        h, c = h
        if submess.numel() == 0 or self.depth == 0:
            if submess.numel() == 0:
                return h * 1.0, c * 1.0
            z = h.new_zeros(submess.numel(), h.shape[1])
            return Fn.index_scatter(z, h, submess), Fn.index_scatter(z, c, submess)
        sub, nbr, K = self._sparse_args(h, fmess, submess, bgraph)
        if c.shape != h.shape:
            raise ValueError("h and c tables differ in shape")
        Wx, T, U4 = self._packed_for_steps(packing.pack_lstm)
        return Fn.SparseLSTM.apply(h, c, packing.feature_columns(fmess, self.input_size), Wx, T, U4, sub, nbr, K, self.depth)
