"""Property heads behind the latent head — drop-in for ggpm/property_optimizer.py (PropertyOptimizer :5-51,
PropertyRegressor :54-66): two independent MLP regressors (HOMO, LUMO) on the two halves of the latent
vector (ggpm/property_vae.py:225-227), each Linear(In, Hd) + ReLU + Dropout [+ ...] + Linear(Hd_last, 1), with a
mean-squared-error loss per head.

Same constructor arguments, attribute names, parameter names / shapes / registration order (`homo_linear.linear.0.weight`,
... — state_dicts interchange with the reference's), same `forward` / `predict` / `compute_loss` contract.  The shipped
shape — one hidden layer (`linear_hidden_size` is an int in every config) — runs both heads, forward and backward, as ONE
kernel launch per direction (csrc/heads.cu); `nn.Dropout` stays a torch op so the RNG stream is the reference's: the
multiplier it would apply is drawn with the same call on a same-shaped tensor, in the same order (HOMO, then LUMO), and
handed to the kernel.  Deeper regressors (a list of hidden sizes) fall back to the library's fused linears layer by layer.
"""
from __future__ import annotations

import ctypes

import torch
from torch import nn

from . import functional as Fn
from ._lib import ACT_RELU, call, fptr, stream_ptr

_MAX_IN, _MAX_W1 = 64, 8192          # limits of the fused kernel (csrc/heads.cu)


class PropertyRegressor(nn.Module):
    def __init__(self, hidden_size: list, dropout: float):
        super(PropertyRegressor, self).__init__()
        self.linear = nn.ModuleList()
        for idx in range(len(hidden_size) - 1):
            self.linear.extend([nn.Linear(hidden_size[idx], hidden_size[idx + 1]),
                                nn.ReLU(), nn.Dropout(dropout)])
        self.linear.append(nn.Linear(hidden_size[-1], 1))

    def forward(self, x):
        """[B, In] -> [B, 1] (ggpm/property_optimizer.py:63-66), layer by layer on the library's linears"""
        squeeze = x.dim() == 1
        h = x.unsqueeze(0) if squeeze else x
        n = len(self.linear)
        for i in range(0, n - 1, 3):
            lin, drop = self.linear[i], self.linear[i + 2]
            h = drop(Fn.linear_act_split([h], lin.weight, [lin.in_features], lin.bias, act=ACT_RELU)[:, :lin.out_features])
        last = self.linear[n - 1]
        y = Fn.linear_act_split([h], last.weight, [last.in_features], last.bias)[:, :1]
        return y[0] if squeeze else y


class _HeadPair(torch.autograd.Function):
    """both one-hidden-layer heads in one launch: (xh, xl, th, tl, mh, ml, 8 parameters) -> (yh [B], yl [B], lossh, lossl)"""

    @staticmethod
    def forward(ctx, xh, xl, th, tl, mh, ml, W1h, b1h, w2h, b2h, W1l, b1l, w2l, b2l):
        B, In = xh.shape
        Hd = W1h.shape[0]
        dev = xh.device
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        hid, out, sq, loss = f(2, B, Hd), f(2, B), f(2, B), torch.zeros(2, dtype=torch.float32, device=dev)
        xs = [xh if xh.stride(1) == 1 else xh.contiguous(), xl if xl.stride(1) == 1 else xl.contiguous()]
        par = [[t.contiguous() for t in (W1h, b1h, w2h.reshape(-1), b2h)], [t.contiguous() for t in (W1l, b1l, w2l.reshape(-1), b2l)]]
        tg = [None if t is None else t.to(torch.float32).reshape(-1).contiguous() for t in (th, tl)]
        ms = [None if m is None else m.contiguous() for m in (mh, ml)]
        ctx.keep = (xs, par, tg, ms, hid, out)
        ctx.dims = (B, In, Hd)
        _HeadPair._call("ggpm_mlp_heads_fwd", ctx, sq, loss, None, None, None, None)
        return out[0], out[1], loss[0], loss[1]

    @staticmethod
    def _call(name, ctx, sq, loss, douts, dlosses, dxs, dpars, db2=None, ld_dx=0):
        xs, par, tg, ms, hid, out = ctx.keep
        B, In, Hd = ctx.dims
        ptrs = []
        P = lambda t: None if t is None else t.data_ptr()
        for h in range(2):
            ptrs += [P(xs[h]), P(par[h][0]), P(par[h][1]), P(par[h][2]), P(par[h][3]), P(ms[h]), P(tg[h]), P(hid[h]), P(out[h]),
                     None if sq is None else P(sq[h]), None if loss is None else P(loss[h:h + 1]),
                     None if douts is None else P(douts[h]), None if dlosses is None else P(dlosses[h]),
                     None if dxs is None else P(dxs[h]),
                     None if dpars is None else P(dpars[h][0]), None if dpars is None else P(dpars[h][1]),
                     None if dpars is None else P(dpars[h][2])]
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        ld_m = Hd if ms[0] is None else ms[0].stride(0)
        st = stream_ptr(xs[0].device)
        if name == "ggpm_mlp_heads_fwd":
            call(name, arr, xs[0].stride(0), xs[1].stride(0), ld_m, B, In, Hd, st)
        else:
            call(name, arr, fptr(db2), xs[0].stride(0), xs[1].stride(0), ld_m, ld_dx, B, In, Hd, st)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dyh, dyl, dlh, dll):
        xs, par, tg, ms, hid, out = ctx.keep
        B, In, Hd = ctx.dims
        dev = out.device
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        douts = [None if d is None else d.contiguous() for d in (dyh, dyl)]
        dlosses = [None if d is None else d.reshape(1).contiguous() for d in (dlh, dll)]
        dxs = [f(B, In), f(B, In)]
        dpars = [[f(Hd, In), f(Hd), f(Hd)], [f(Hd, In), f(Hd), f(Hd)]]
        db2 = f(2)
        _HeadPair._call("ggpm_mlp_heads_bwd", ctx, None, None, douts, dlosses, dxs, dpars, db2, In)
        gh, gl = dpars
        return (dxs[0], dxs[1], None, None, None, None, gh[0], gh[1], gh[2].reshape(1, Hd), db2[0:1], gl[0], gl[1],
                gl[2].reshape(1, Hd), db2[1:2])


def _flatten(x):
    """ggpm/property_optimizer.py:40-46: drop a trailing singleton dimension"""
    if x.shape[-1] == 1:
        if len(x.shape) == 1:
            x = x[-1]
        elif len(x.shape) > 1:
            x = x[:, -1]
    return x


class PropertyOptimizer(nn.Module):
    def __init__(self, input_size, hidden_size, dropout):
        super(PropertyOptimizer, self).__init__()
        self.input_size = input_size
        hidden_size = [hidden_size] if isinstance(hidden_size, int) else hidden_size
        hidden_size = [input_size] + list(hidden_size)
        self.homo_linear = PropertyRegressor(hidden_size, dropout)
        self.lumo_linear = PropertyRegressor(hidden_size, dropout)

    def compute_loss(self, outputs, targets):
        return torch.nn.MSELoss(reduction='mean')(outputs, targets)

    # -- fused path --------------------------------------------------------------------------------------------------
    def _fusable(self, homo_vecs, lumo_vecs):
        a, b = self.homo_linear.linear, self.lumo_linear.linear
        if len(a) != 4 or len(b) != 4 or homo_vecs.dim() != 2 or lumo_vecs.dim() != 2 or homo_vecs.shape != lumo_vecs.shape:
            return False
        In, Hd = a[0].in_features, a[0].out_features
        return (homo_vecs.dtype == torch.float32 and homo_vecs.is_cuda == lumo_vecs.is_cuda and In <= _MAX_IN and Hd * In <= _MAX_W1
                and b[0].in_features == In and b[0].out_features == Hd and homo_vecs.shape[1] == In)

    def _dropout_multiplier(self, drop: nn.Dropout, B, Hd, like):
        """what drop(x) multiplies x [B, Hd] by: the same F.dropout call on a same-shaped tensor consumes the generator
        exactly as the reference's nn.Dropout does (ggpm/property_optimizer.py:58-59)"""
        if not drop.training or drop.p == 0.0:
            return None
        return drop(torch.ones(B, Hd, dtype=like.dtype, device=like.device))

    def _heads(self, homo_vecs, lumo_vecs, targets):
        a, b = self.homo_linear.linear, self.lumo_linear.linear
        B, Hd = homo_vecs.shape[0], a[0].out_features
        mh = self._dropout_multiplier(a[2], B, Hd, homo_vecs)            # HOMO's draw first, as the reference's call order
        ml = self._dropout_multiplier(b[2], B, Hd, lumo_vecs)
        th, tl = (None, None) if targets is None else targets
        return _HeadPair.apply(homo_vecs, lumo_vecs, th, tl, mh, ml, a[0].weight, a[0].bias, a[3].weight, a[3].bias,
                               b[0].weight, b[0].bias, b[3].weight, b[3].bias)

    # -- reference interface -----------------------------------------------------------------------------------------
    def forward(self, homo_vecs, lumo_vecs, targets):
        """-> (homo_loss, lumo_loss, homo_outputs, lumo_outputs), ggpm/property_optimizer.py:25-37"""
        if self._fusable(homo_vecs, lumo_vecs) and all(torch.is_tensor(t) and t.numel() == homo_vecs.shape[0] for t in targets):
            yh, yl, lh, ll = self._heads(homo_vecs, lumo_vecs, targets)
            return lh, ll, yh, yl
        homo_outputs, lumo_outputs = self.predict(homo_vecs, lumo_vecs)
        return (self.compute_loss(homo_outputs, targets[0]), self.compute_loss(lumo_outputs, targets[1]),
                homo_outputs, lumo_outputs)

    def predict(self, homo_vecs, lumo_vecs):
        if self._fusable(homo_vecs, lumo_vecs):
            yh, yl, _, _ = self._heads(homo_vecs, lumo_vecs, None)
            return yh, yl
        return _flatten(self.homo_linear(homo_vecs)), _flatten(self.lumo_linear(lumo_vecs))
