"""Latent head (ggpm_b200/latent.py, csrc/latent.cu): oracle vs the reference-made golden vectors
and vs the live reference; host logic on the C-ABI emulator (CPU); the CUDA kernels through the C ABI (GPU)."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import ROOT                                                              # noqa: E402
import _emu_lib                                                                        # noqa: E402
from ggpm_b200 import _lib, latent, nnutils                                            # noqa: E402
from oracle import latent_oracle as lo                                                 # noqa: E402
from oracle.ref_loader import load_reference, reference_available                       # noqa: E402

Z = np.load(os.path.join(ROOT, "tests", "golden_latent", "latent_rsample.npz"))
CASES = [(c, p) for c in ("a", "b", "c") for p in (1, 0)]
TOL = 1e-5          # fp32 path vs the fp64 reference run


def _rel(a, b):
    a = a.detach().cpu().double().numpy() if isinstance(a, torch.Tensor) else np.asarray(a, dtype=np.float64)
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-30)


def _inputs(case, dtype, device="cpu"):
    t = lambda k: torch.from_numpy(Z[case + "_" + k]).to(dtype).to(device)
    return t("in_z"), t("wm"), t("bm"), t("wv"), t("bv"), t("eps"), t("cot")


@pytest.mark.parametrize("case,perturb", CASES)
def test_oracle_matches_reference_golden(case, perturb):
    z, wm, bm, wv, bv, eps, cot = _inputs(case, torch.float64)
    leaves = [x.clone().requires_grad_(True) for x in (z, wm, bm, wv, bv)]
    zs, kl = lo.rsample(*leaves, eps, bool(perturb))
    ((zs * cot).sum() + 0.7 * kl).backward()
    tag = "%s_%d_" % (case, perturb)
    assert _rel(zs, Z[tag + "z"]) <= 1e-12 and _rel(kl, Z[tag + "kl"]) <= 1e-12
    for leaf, k in zip(leaves, ("dz_in", "dwm", "dbm", "dwv", "dbv")):
        assert _rel(leaf.grad, Z[tag + k]) <= 1e-12, k


@pytest.mark.skipif(not reference_available(), reason="reference checkout not present")
def test_oracle_matches_live_reference_and_its_noise_draw():
    load_reference()
    hpgnn = importlib.import_module("ggpm.hpgnn")
    torch.manual_seed(5)
    Wm, Wv, z = torch.nn.Linear(40, 6), torch.nn.Linear(40, 6), torch.randn(9, 40)
    for perturb in (True, False):
        torch.manual_seed(77)
        eps = torch.randn(9, 6)
        torch.manual_seed(77)
        zs, kl = hpgnn.HierVAE.rsample(None, z, Wm, Wv, perturb)
        after_ref = torch.rand(1)
        zo, klo = lo.rsample(z, Wm.weight, Wm.bias, Wv.weight, Wv.bias, eps, perturb)
        assert torch.allclose(zs, zo, atol=1e-6) and torch.allclose(kl, klo, atol=1e-6)
        # the product draws the same noise from the same generator state, and leaves the generator where the reference does
        _lib._install_test_backend(_emu_lib.EmuLib())
        nnutils.set_default_device("cpu")
        try:
            torch.manual_seed(77)
            zp, klp = latent.rsample(z, Wm, Wv, perturb)
            assert torch.equal(torch.rand(1), after_ref)
        finally:
            _lib._install_test_backend(None)
            nnutils.set_default_device(None)
        assert torch.allclose(zp, zs, atol=1e-5) and torch.allclose(klp, kl, atol=1e-5)


def _check_product(case, perturb, device):
    z, wm, bm, wv, bv, eps, cot = _inputs(case, torch.float32, device)
    H, L = wm.shape[1], wm.shape[0]
    Wm, Wv = torch.nn.Linear(H, L).to(device), torch.nn.Linear(H, L).to(device)
    with torch.no_grad():
        Wm.weight.copy_(wm); Wm.bias.copy_(bm); Wv.weight.copy_(wv); Wv.bias.copy_(bv)
    zin = z.clone().requires_grad_(True)
    zs, kl = latent.rsample(zin, Wm, Wv, bool(perturb), epsilon=eps)
    assert zs.shape == (z.shape[0], L) and zs.is_contiguous() and kl.dim() == 0
    ((zs * cot).sum() + 0.7 * kl).backward()
    tag = "%s_%d_" % (case, perturb)
    assert _rel(zs, Z[tag + "z"]) <= TOL and _rel(kl, Z[tag + "kl"]) <= TOL
    for g, k in ((zin.grad, "dz_in"), (Wm.weight.grad, "dwm"), (Wm.bias.grad, "dbm"), (Wv.weight.grad, "dwv"),
                 (Wv.bias.grad, "dbv")):
        assert _rel(g, Z[tag + k]) <= 2 * TOL, k


@pytest.fixture()
def emu():
    lib = _emu_lib.EmuLib()
    _lib._install_test_backend(lib)
    nnutils.set_default_device("cpu")
    yield lib
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)


@pytest.mark.parametrize("case,perturb", CASES)
def test_host_logic_on_emulator_matches_reference_golden(emu, case, perturb):
    _check_product(case, perturb, "cpu")
    assert "rsample_fwd" in emu.calls and "rsample_bwd" in emu.calls


def test_kl_only_and_z_only_gradients(emu):
    z, wm, bm, wv, bv, eps, cot = _inputs("b", torch.float32)
    Wm, Wv = torch.nn.Linear(48, 5), torch.nn.Linear(48, 5)
    for which in ("kl", "z"):
        Wm.zero_grad(); Wv.zero_grad()
        zs, kl = latent.rsample(z, Wm, Wv, True, epsilon=eps)
        (kl if which == "kl" else (zs * cot).sum()).backward()
        zo, klo = lo.rsample(z.double(), *[p.detach().double().requires_grad_(True) for p in (Wm.weight, Wm.bias, Wv.weight, Wv.bias)],
                             eps.double(), True)
        assert _rel(zs, zo.detach().numpy()) <= TOL and Wm.weight.grad is not None and Wv.weight.grad is not None
    with pytest.raises(ValueError):
        latent.rsample(z[:, :10], Wm, Wv)


@pytest.mark.gpu
@pytest.mark.parametrize("case,perturb", CASES)
def test_cuda_kernels_match_reference_golden(case, perturb):
    _check_product(case, perturb, "cuda")


@pytest.mark.gpu
def test_cuda_repeatable_and_noise_comes_from_the_device_generator():
    torch.manual_seed(3)
    Wm, Wv = torch.nn.Linear(250, 24).cuda(), torch.nn.Linear(250, 24).cuda()
    z = torch.randn(256, 250, device="cuda")
    outs = []
    for _ in range(3):
        torch.manual_seed(11)
        outs.append(latent.rsample(z, Wm, Wv, True))
    # same seed -> same noise; the projection is summed from K-split partial sums (fp32 reductions), so repeats agree
    # to rounding, not bit for bit
    assert all(_rel(o[0], outs[0][0].detach().double().cpu().numpy()) <= 1e-6 and _rel(o[1], outs[0][1].detach().double().cpu().numpy()) <= 1e-6
               for o in outs[1:])
    torch.manual_seed(11)
    eps = torch.randn(256, 24, device="cuda")
    zo, klo = lo.rsample(z.double().cpu(), *[p.detach().double().cpu() for p in (Wm.weight, Wm.bias, Wv.weight, Wv.bias)],
                         eps.double().cpu(), True)
    assert _rel(outs[0][0], zo.numpy()) <= TOL and _rel(outs[0][1], klo.numpy()) <= TOL
