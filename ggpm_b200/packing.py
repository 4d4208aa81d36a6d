"""Gate-weight packing for the CUDA message-passing kernels (differentiable torch reshuffling).

The reference stores every gate as nn.Linear(IN + H -> H) applied to the concatenation [x | h]
(ggpm/rnn.py:13-16, 69-72), with x = [node feature of the source (F columns) | one-hot columns]
(ggpm/encoder.py:103-106, 114-116, 121-125).  The kernels want, per layer:

  Wx  [G*Hp, Fp]    x-columns that multiply the node feature, gates stacked along rows (Fp = F rounded up to 4)
  T   [ntab, G*Hp]  the one-hot columns transposed into lookup rows; rows [0, n_bias_rows) (the ones
                    indexed by k1, exactly one of which every message selects) also carry the biases
  U*  [.., Hp]      the h-columns, zero padded from H to Hp = multiple of 32

These are built with ordinary autograd-tracked torch ops on the nn.Parameters, so parameter
gradients flow back through the packing without hand-written unpacking, the module keeps the
reference's state_dict, and nothing is cached across optimizer steps.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F_


def feature_width(F: int) -> int:
    """K of the hoisted x-projection GEMM: F rounded up to 4 floats (16-byte rows for the tensor-core path)"""
    return (F + 3) // 4 * 4


def feature_columns(phi: torch.Tensor, F: int) -> torch.Tensor:
    """node-feature table at the width Wx is packed for: [N1, feature_width(F)].  Columns >= F multiply zero
    weight columns; a table that is already wider is used in place (a view), a narrower one is zero-padded."""
    Fp = feature_width(F)
    if phi.shape[1] >= Fp:
        out = phi[:, :Fp]
    else:
        out = F_.pad(phi[:, :F], (0, Fp - F))
    if out.stride(1) != 1 or (out.shape[0] > 1 and out.stride(0) % 4 != 0) or out.data_ptr() % 16 != 0:
        out = out.contiguous()
    return out


def _pack_common(weights, biases, F, IN, H, Hp, n_bias_rows):
    """weights: list of [H, IN(+H)] gate matrices; biases: list of [H] or None (same length)."""
    G = len(weights)
    Fp = feature_width(F)
    Wall = torch.stack([w[:, :IN] for w in weights])                        # [G, H, IN]
    Wx = F_.pad(Wall[:, :, :F], (0, Fp - F, 0, Hp - H)).reshape(G * Hp, Fp)  # [G*Hp, Fp], zero pad columns
    ntab = IN - F
    ball = torch.stack([b if b is not None else torch.zeros_like(weights[0][:, 0]) for b in biases])  # [G, H]
    if ntab == 0:
        # dense message inputs (no one-hot columns): a single lookup row that carries the biases
        ntab, T = 1, ball.unsqueeze(0)
    else:
        T = Wall[:, :, F:IN].permute(2, 0, 1)                               # [ntab, G, H]
        carrier = torch.zeros(ntab, 1, 1, dtype=Wall.dtype, device=Wall.device)
        carrier[:n_bias_rows] = 1.0
        T = T + carrier * ball.unsqueeze(0)
    T = F_.pad(T, (0, Hp - H)).reshape(ntab, G * Hp)
    return Wx, T


def _pad_square(u, H, Hp):
    return F_.pad(u, (0, Hp - H, 0, Hp - H))


def _ptr_array(tensors):
    import ctypes
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


class PackLSTM(torch.autograd.Function):
    """(W_i, b_i, W_o, b_o, W, b, W_f, b_f) -> (Wx [4Hp, Fp], T [ntab, 4Hp], U4 [4Hp, Hp]) in ONE launch
    (ggpm_lstm_pack); backward is one launch too (ggpm_lstm_unpack).  Same result as the torch reshuffling of
    _pack_common (kept for the GRU), without its ~30 tiny kernels and autograd nodes per layer and step."""

    @staticmethod
    def forward(ctx, F, IN, H, Hp, n_bias_rows, *wb):
        from ._lib import call, fptr, stream_ptr
        ws = [w.contiguous() for w in wb[0::2]]
        bs = [b.contiguous() for b in wb[1::2]]
        dev = ws[0].device
        Fp = feature_width(F)
        ntab = IN - F if IN > F else 1
        Wx = torch.empty(4 * Hp, Fp, dtype=torch.float32, device=dev)
        T = torch.empty(ntab, 4 * Hp, dtype=torch.float32, device=dev)
        U4 = torch.empty(4 * Hp, Hp, dtype=torch.float32, device=dev)
        for t in ws + bs:
            fptr(t, "gate parameter")
        call("ggpm_lstm_pack", _ptr_array(ws), _ptr_array(bs), H, Hp, F, Fp, IN, ntab, n_bias_rows, fptr(Wx), fptr(T),
             fptr(U4), stream_ptr(dev))
        ctx.dims = (F, IN, H, Hp, n_bias_rows, Fp)
        ctx.dev = dev
        return Wx, T, U4

    @staticmethod
    def backward(ctx, dWx, dT, dU4):
        from ._lib import call, fptr, stream_ptr
        F, IN, H, Hp, n_bias_rows, Fp = ctx.dims
        dev = ctx.dev
        dws = [torch.empty(H, IN + H, dtype=torch.float32, device=dev) for _ in range(4)]
        dbs = [torch.empty(H, dtype=torch.float32, device=dev) for _ in range(4)]
        g = [None if t is None else t.contiguous() for t in (dWx, dT, dU4)]
        call("ggpm_lstm_unpack", _ptr_array(dws), _ptr_array(dbs), H, Hp, F, Fp, IN, n_bias_rows,
             *[None if t is None else fptr(t) for t in g], stream_ptr(dev))
        out = [None] * 5
        for w, b in zip(dws, dbs):
            out += [w, b]
        return tuple(out)


def pack_lstm(rnn, F: int, Hp: int, n_bias_rows: int):
    """rnn: module with W_i, W_o, W_f, W = Sequential(Linear(IN+H -> H), act).  Gate order [i | o | u | f].
    Returns (Wx, T, U4) with U4 = [U_i; U_o; U_u; U_f] in one [4Hp, Hp] buffer, which lets the library run the gate
    and forget projections (and their gradients) as single tensor-core launches."""
    H, IN = rnn.hidden_size, rnn.input_size
    lin = [rnn.W_i[0], rnn.W_o[0], rnn.W[0], rnn.W_f[0]]
    wb = []
    for l in lin:
        wb += [l.weight, l.bias]
    return PackLSTM.apply(F, IN, H, Hp, n_bias_rows, *wb)


def pack_lstm_torch(rnn, F: int, Hp: int, n_bias_rows: int):
    """rnn: module with W_i, W_o, W_f, W = Sequential(Linear(IN+H -> H), act).  Gate order [i | o | u | f]."""
    H, IN = rnn.hidden_size, rnn.input_size
    lin = [rnn.W_i[0], rnn.W_o[0], rnn.W[0], rnn.W_f[0]]
    Wx, T = _pack_common([l.weight for l in lin], [l.bias for l in lin], F, IN, H, Hp, n_bias_rows)
    # one [4Hp, Hp] buffer [U_i; U_o; U_u; U_f]: Ug / Uf are views of it, which lets the library run the gate and
    # forget projections (and their gradients) as single tensor-core launches
    U4 = torch.cat([_pad_square(l.weight[:, IN:], H, Hp) for l in lin], dim=0)
    return Wx, T, U4[:3 * Hp], U4[3 * Hp:]


def pack_gru(rnn, F: int, Hp: int, n_bias_rows: int):
    """rnn: module with W_z, W_h = Linear(IN+H -> H), W_r = Linear(IN -> H, no bias), U_r = Linear(H -> H).
    x-gate order [z | h | r]."""
    H, IN = rnn.hidden_size, rnn.input_size
    Wx, T = _pack_common([rnn.W_z.weight, rnn.W_h.weight, rnn.W_r.weight], [rnn.W_z.bias, rnn.W_h.bias, None],
                         F, IN, H, Hp, n_bias_rows)
    Uz = _pad_square(rnn.W_z.weight[:, IN:], H, Hp)
    Uh = _pad_square(rnn.W_h.weight[:, IN:], H, Hp)
    Ur = _pad_square(rnn.U_r.weight, H, Hp)
    br = F_.pad(rnn.U_r.bias, (0, Hp - H))
    return Wx, T, Uz, Ur, br, Uh
