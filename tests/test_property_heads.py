"""Property heads (ggpm_b200/property_optimizer.py, csrc/heads.cu): the oracle against the
reference-made golden vectors and the live reference; the drop-in classes on the C-ABI emulator (CPU) and through the CUDA
kernels (GPU): outputs, both MSE losses, input gradients and every parameter gradient, with the reference's own dropout
multipliers; state_dict interchange with the reference class; the RNG stream of train-mode dropout."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import ROOT                                                              # noqa: E402
import _checks as C                                                                     # noqa: E402
import _emu_lib                                                                        # noqa: E402
import ggpm_b200                                                                       # noqa: E402
from ggpm_b200 import _lib, nnutils                                                     # noqa: E402
from oracle import property_oracle as po                                               # noqa: E402
from oracle.ref_loader import load_reference_module, reference_available                # noqa: E402

Z = np.load(os.path.join(ROOT, "tests", "golden_property", "property_heads.npz"))
CASES = {"a": (256, 12, 128, 0.0), "b": (37, 12, 128, 0.25), "c": (20, 8, [48], 0.25), "d": (9, 6, [16, 10], 0.0)}
TOL = 1e-5


def _sd(case, dtype):
    pre = case + "_sd_"
    return {k[len(pre):]: torch.from_numpy(Z[k]).to(dtype) for k in Z.files if k.startswith(pre)}


def _drop(case):
    pre = case + "_drop_"
    return {k[len(pre):]: torch.from_numpy(Z[k]) for k in Z.files if k.startswith(pre)}


def _t(case, k, dtype, device="cpu"):
    return torch.from_numpy(Z[case + "_" + k]).to(dtype).to(device)


@pytest.mark.parametrize("case", sorted(CASES))
def test_oracle_matches_reference_golden(case):
    leaves = {k: v.requires_grad_(True) for k, v in _sd(case, torch.float64).items()}
    a, b = _t(case, "xh", torch.float64).requires_grad_(True), _t(case, "xl", torch.float64).requires_grad_(True)
    lh, ll, yh, yl = po.forward(leaves, a, b, (_t(case, "th", torch.float64), _t(case, "tl", torch.float64)), _drop(case))
    (1.3 * lh + 0.7 * ll + (yh * _t(case, "coth", torch.float64)).sum() + (yl * _t(case, "cotl", torch.float64)).sum()).backward()
    for got, key in ((lh, "lh"), (ll, "ll"), (yh, "yh"), (yl, "yl"), (a.grad, "dxh"), (b.grad, "dxl")):
        assert C.rel(got, Z[case + "_" + key]) <= 1e-12, key
    for k, v in leaves.items():
        assert C.rel(v.grad, Z[case + "_g_" + k]) <= 1e-12, k


@pytest.mark.skipif(not reference_available(), reason="no copy of the reference reachable")
def test_state_dict_interchanges_with_the_reference_class():
    ref = load_reference_module("property_optimizer")
    for hidden in (128, [32, 16]):
        theirs, ours = ref.PropertyOptimizer(12, hidden, 0.1), ggpm_b200.PropertyOptimizer(12, hidden, 0.1)
        assert list(theirs.state_dict().keys()) == list(ours.state_dict().keys())
        assert [tuple(v.shape) for v in theirs.state_dict().values()] == [tuple(v.shape) for v in ours.state_dict().values()]
        assert [n for n, _ in theirs.named_parameters()] == [n for n, _ in ours.named_parameters()]
        ours.load_state_dict(theirs.state_dict(), strict=True)
        theirs.load_state_dict(ours.state_dict(), strict=True)
        assert ours.input_size == theirs.input_size


def _check_product(case, device):
    B, In, hidden, p = CASES[case]
    model = ggpm_b200.PropertyOptimizer(In, hidden, p)
    model.load_state_dict(_sd(case, torch.float32), strict=True)
    model = model.to(device).train()
    drop = _drop(case)
    for prefix, reg in (("homo_linear.", model.homo_linear), ("lumo_linear.", model.lumo_linear)):
        for i, layer in enumerate(reg.linear):
            if isinstance(layer, torch.nn.Dropout) and (prefix + str(i)) in drop:
                reg.linear[i] = C.FixedDrop(drop[prefix + str(i)].float())
    a = _t(case, "xh", torch.float32, device).requires_grad_(True)
    b = _t(case, "xl", torch.float32, device).requires_grad_(True)
    lh, ll, yh, yl = model(a, b, (_t(case, "th", torch.float32, device), _t(case, "tl", torch.float32, device)))
    assert yh.shape == (B,) and yl.shape == (B,) and lh.dim() == 0
    (1.3 * lh + 0.7 * ll + (yh * _t(case, "coth", torch.float32, device)).sum() + (yl * _t(case, "cotl", torch.float32, device)).sum()).backward()
    tag = "property_heads:" + case
    for got, key in ((lh, "lh"), (ll, "ll"), (yh, "yh"), (yl, "yl"), (a.grad, "dxh"), (b.grad, "dxl")):
        C.record(tag, key, C.rel(got, Z[case + "_" + key]), TOL)
        assert C.rel(got, Z[case + "_" + key]) <= TOL, (key, C.rel(got, Z[case + "_" + key]))
    for k, q in model.named_parameters():
        C.record(tag, "grad:" + k, C.rel(q.grad, Z[case + "_g_" + k]), TOL)
        assert C.rel(q.grad, Z[case + "_g_" + k]) <= TOL, (k, C.rel(q.grad, Z[case + "_g_" + k]))
    # predict(): same outputs, no targets
    ph, pl = model.predict(a.detach(), b.detach())
    assert C.rel(ph, Z[case + "_yh"]) <= TOL and C.rel(pl, Z[case + "_yl"]) <= TOL
    return model


@pytest.fixture()
def emu():
    lib = _emu_lib.EmuLib()
    _lib._install_test_backend(lib)
    nnutils.set_default_device("cpu")
    yield lib
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)


@pytest.mark.parametrize("case", sorted(CASES))
def test_host_logic_on_emulator_matches_reference_golden(emu, case):
    _check_product(case, "cpu")
    fused = case in ("a", "b", "c")
    assert ("mlp_heads_fwd" in emu.calls) == fused and ("mlp_heads_bwd" in emu.calls) == fused


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_cuda_kernels_match_reference_golden(case):
    before = _lib.launch_count()
    _check_product(case, "cuda")
    assert _lib.launch_count() > before


@pytest.mark.gpu
def test_train_mode_dropout_draws_what_nn_dropout_would():
    """same seed, same device: the fused heads must see exactly the multipliers a plain torch evaluation of the same
    modules (Linear, ReLU, nn.Dropout, Linear — the reference's layer list) draws, HOMO first, then LUMO"""
    torch.manual_seed(3)
    model = ggpm_b200.PropertyOptimizer(12, 128, 0.3).cuda().train()
    xh, xl = torch.randn(64, 12, device="cuda"), torch.randn(64, 12, device="cuda")
    torch.manual_seed(21)
    yh, yl = model.predict(xh, xl)
    torch.manual_seed(21)
    ref = []
    for reg, x in ((model.homo_linear, xh), (model.lumo_linear, xl)):
        h = x
        for layer in reg.linear:                       # plain torch, as ggpm/property_optimizer.py:63-66
            h = layer(h)
        ref.append(h[:, -1])
    assert C.rel(yh, ref[0]) <= TOL and C.rel(yl, ref[1]) <= TOL
