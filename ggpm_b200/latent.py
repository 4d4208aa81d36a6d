"""Latent head on the encoder output — drop-in for ``PropertyVAE.rsample`` /
``HierVAE.rsample`` (ggpm/property_vae.py:26-34): same arguments, same return values ``(z_vecs, kl_loss)``, same draw
from the global torch generator (``torch.randn_like(z_mean)``, drawn whether or not it is used, as the reference does),
so a model that swaps it in keeps its RNG stream.

The two ``nn.Linear(hidden, latent)`` projections run as ONE fused linear over ``[W_mean; W_var]`` on the library's GEMM
path; the reparameterisation, the KL term and their backward are one hand-written kernel each (csrc/latent.cu).
"""
from __future__ import annotations

import torch

from . import functional as Fn
from ._lib import call, fptr, stream_ptr


class _RSample(torch.autograd.Function):
    """(Y [B, >= 2L] = [z_mean | v | pad], eps [B, L] or None) -> (z [B, L], kl scalar)"""

    @staticmethod
    def forward(ctx, Y, eps, L):
        B, width = Y.shape
        z = torch.empty(B, L, dtype=torch.float32, device=Y.device)
        kl = torch.empty((), dtype=torch.float32, device=Y.device)
        row_kl = torch.empty(B, dtype=torch.float32, device=Y.device)
        call("ggpm_rsample_fwd", fptr(Y), Y.stride(0), L, None if eps is None else fptr(eps), 0 if eps is None else eps.stride(0),
             B, L, fptr(z), L, fptr(kl), fptr(row_kl), stream_ptr(Y.device))
        ctx.save_for_backward(Y, eps)
        ctx.L = L
        return z, kl

    @staticmethod
    def backward(ctx, dz, dkl):
        Y, eps = ctx.saved_tensors
        B, width = Y.shape
        dY = torch.empty(B, width, dtype=torch.float32, device=Y.device)
        dz = None if dz is None else dz.contiguous()
        dkl = None if dkl is None else dkl.contiguous()
        call("ggpm_rsample_bwd", fptr(Y), Y.stride(0), ctx.L, None if eps is None else fptr(eps),
             0 if eps is None else eps.stride(0), B, ctx.L, None if dz is None else fptr(dz), 0 if dz is None else dz.stride(0),
             None if dkl is None else fptr(dkl), fptr(dY), width, width, stream_ptr(Y.device))
        return dY, None, None


def rsample(z_vecs: torch.Tensor, W_mean: torch.nn.Linear, W_var: torch.nn.Linear, perturb: bool = True, epsilon=None):
    """``epsilon`` (optional, [B, L]) overrides the draw — used by the parity tests; the reference has no such argument."""
    if W_mean.out_features != W_var.out_features or W_mean.in_features != W_var.in_features:
        raise ValueError("W_mean and W_var must have the same shape")
    L, H = W_mean.out_features, W_mean.in_features
    if z_vecs.dim() != 2 or z_vecs.shape[1] != H:
        raise ValueError("z_vecs must be [batch, %d]" % H)
    W = torch.cat([W_mean.weight, W_var.weight], dim=0)
    b = None if W_mean.bias is None else torch.cat([W_mean.bias, W_var.bias], dim=0)
    Y = Fn.linear_act_split([z_vecs], W, [H], b)                    # [B, round_up(2L, 32)]: [z_mean | v | zeros]
    if epsilon is None:
        epsilon = torch.randn(z_vecs.shape[0], L, dtype=z_vecs.dtype, device=z_vecs.device)   # randn_like(z_mean), :32
    eps = epsilon.to(torch.float32).contiguous() if perturb else None
    return _RSample.apply(Y, eps, L)
