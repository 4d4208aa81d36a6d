"""B200: the allocation-free entry point `ggpm_rsample_fwd_noalloc` (include/ggpm_b200.h; no counterpart in the
reference) against the existing `ggpm_rsample_fwd`, both through the C-ABI, on two calls in a row.

The variant runs the same two kernels with the per-row KL scratch in a file-scope `static __device__` array (65 536
floats) instead of a caller buffer, so the results must be IDENTICAL, not merely close; the test still states the
suite's tolerance (1e-5 of max|ref|, tests/_checks.py) and additionally asserts bit equality.  The second call uses
fewer rows and different data than the first, so stale contents of the shared array beyond its B rows, or left by the
first call, would show.  Collected last on purpose.
"""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _checks as C                                                                     # noqa: E402
from ggpm_b200 import _lib                                                              # noqa: E402

pytestmark = pytest.mark.gpu
MAX_ROWS = 65536                      # kRowKlMax of ggpm_b200/csrc/latent.cu


def _inputs(seed, B, L, width, with_eps=True):
    gen = torch.Generator().manual_seed(seed)
    Y = torch.randn(B, width, generator=gen).cuda()
    eps = torch.randn(B, L, generator=gen).cuda() if with_eps else None
    return Y, eps


def _run(name, Y, eps, L, var_off, with_scratch):
    B = Y.shape[0]
    z = torch.full((B, L), float("nan"), device="cuda")
    kl = torch.full((), float("nan"), device="cuda")
    args = [_lib.fptr(Y), Y.stride(0), var_off, None if eps is None else _lib.fptr(eps), 0 if eps is None else eps.stride(0),
            B, L, _lib.fptr(z), L, _lib.fptr(kl)]
    keep = None
    if with_scratch:
        keep = torch.empty(B, device="cuda")
        args.append(_lib.fptr(keep))
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    _lib.call(name, *args, _lib.stream_ptr(Y.device))
    after = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    torch.cuda.synchronize()
    return z, kl, before == after


def test_allocation_free_rsample_matches_the_existing_entry_point_on_two_calls_in_a_row():
    assert _lib._TEST_BACKEND is None
    launches = _lib.launch_count()
    # (seed, rows, latent, row width, offset of the variance block, perturb)
    for call_no, (seed, B, L, width, var_off, with_eps) in enumerate([(1, 4096, 24, 64, 32, True), (2, 1000, 24, 48, 24, True),
                                                                      (3, 7, 5, 16, 8, False)]):
        Y, eps = _inputs(seed, B, L, width, with_eps)
        z_ref, kl_ref, _ = _run("ggpm_rsample_fwd", Y, eps, L, var_off, True)
        z, kl, no_torch_alloc = _run("ggpm_rsample_fwd_noalloc", Y, eps, L, var_off, False)
        assert no_torch_alloc
        assert bool(torch.isfinite(z).all()) and bool(torch.isfinite(kl))
        assert C.rel(z, z_ref) <= C.TOL_OUT and C.rel(kl, kl_ref) <= C.TOL_OUT, (call_no, C.rel(z, z_ref), C.rel(kl, kl_ref))
        assert torch.equal(z, z_ref) and torch.equal(kl, kl_ref), call_no          # same kernels, same order of additions
        # and against the formula itself (ggpm/property_vae.py:26-34), fp64 on the host
        y = Y.double().cpu()
        m, lv = y[:, :L], -y[:, var_off:var_off + L].abs()
        kl64 = -0.5 * torch.sum(1.0 + lv - m * m - torch.exp(lv)) / B
        z64 = m + torch.exp(lv / 2) * eps.double().cpu() if with_eps else m
        assert C.rel(z, z64) <= C.TOL_OUT and C.rel(kl, kl64) <= C.TOL_OUT
    assert _lib.launch_count() == launches + 3 * 4


def test_allocation_free_rsample_at_its_maximum_and_beyond():
    L, width = 8, 16
    Y, eps = _inputs(4, MAX_ROWS + 1, L, width)
    z_ref, kl_ref, _ = _run("ggpm_rsample_fwd", Y[:MAX_ROWS], eps[:MAX_ROWS], L, 8, True)
    z, kl, _ = _run("ggpm_rsample_fwd_noalloc", Y[:MAX_ROWS], eps[:MAX_ROWS], L, 8, False)
    assert torch.equal(z, z_ref) and torch.equal(kl, kl_ref)
    with pytest.raises(RuntimeError, match="exceed the built-in scratch"):
        _run("ggpm_rsample_fwd_noalloc", Y, eps, L, 8, False)
    # the refusal leaves the library usable
    z2, kl2, _ = _run("ggpm_rsample_fwd_noalloc", Y[:100], eps[:100], L, 8, False)
    z2_ref, kl2_ref, _ = _run("ggpm_rsample_fwd", Y[:100], eps[:100], L, 8, True)
    assert torch.equal(z2, z2_ref) and torch.equal(kl2, kl2_ref)
