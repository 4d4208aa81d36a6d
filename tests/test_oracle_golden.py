"""CPU: the oracle against the committed golden vectors (made by the unmodified reference,
oracle/make_golden.py) and, when /root/reference exists, against the live reference."""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden
from oracle import encoder_oracle as eo
from oracle.ref_loader import FakePairVocab, IdAtomVocab, load_reference, reference_available

OUT_NAMES = {"hier": ("hroot", "hnode", "hinter", "hatom"), "motif": ("hroot", "hnode")}


def _oracle_run(meta, tree, graph, dtype):
    nm, na = meta["vocab"]
    sd = eo.random_state(meta["kind"], meta["rnn"], meta["hidden"], meta["hidden"], meta["weight_seed"],
                         n_motif=nm, n_attach=na)
    sd = {k: v.to(dtype).requires_grad_(True) for k, v in sd.items()}
    if meta["kind"] == "hier":
        outs = eo.hier_encoder(sd, meta["rnn"], meta["depthT"], meta["depthG"], tree, graph)
    else:
        outs = eo.motif_encoder(sd, meta["rnn"], meta["depthT"], tree)
    return sd, outs


def _rel(a, b):
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-300)


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_golden_fp64(name):
    meta, z, tree, graph = load_golden(name)
    if meta["hidden"] > 64:
        pytest.skip("big case covered by test_oracle_cfg1_outputs")
    sd, outs = _oracle_run(meta, tree, graph, torch.float64)
    names = OUT_NAMES[meta["kind"]]
    for nme, o in zip(names, outs):
        assert _rel(o.detach().numpy(), z["out_" + nme]) <= 1e-12, nme
    sum((o * torch.from_numpy(z["cot_" + n])).sum() for n, o in zip(names, outs)).backward()
    for k, p in sd.items():
        ref = z["gall_" + k]
        got = p.grad.numpy() if p.grad is not None else np.zeros_like(ref)
        if np.abs(ref).max() > 0:
            assert _rel(got, ref) <= 1e-11, k
        else:
            assert np.abs(got).max() == 0
    # rows 0 of the node tables are exactly zero (ggpm/encoder.py:36-38)
    for o in outs[1:]:
        assert float(o[0].abs().max()) == 0.0


def test_oracle_cfg1_outputs():
    """BASELINE config #1 stand-in (H=250, depth 20/20, 20 molecules): forward + hroot.sum() grads."""
    meta, z, tree, graph = load_golden("hier_lstm_cfg1")
    sd, outs = _oracle_run(meta, tree, graph, torch.float64)
    for nme, o in zip(OUT_NAMES["hier"], outs):
        assert _rel(o.detach().numpy(), z["out_" + nme].astype(np.float64)) <= 2e-7, nme   # stored as fp32
    outs[0].sum().backward()
    stride = meta["grad_stride"]
    for k, p in sd.items():
        if "groot_" + k not in z.files:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k     # tree_encoder.W_o
            continue
        assert _rel(p.grad.reshape(-1)[::stride].numpy(), z["groot_" + k]) <= 1e-10, k
        assert abs(float(p.grad.norm()) - float(z["gnorm_" + k])) <= 1e-10 * max(1.0, float(z["gnorm_" + k]))


@pytest.mark.parametrize("name", ["hier_lstm_deep", "hier_gru_small"])
def test_checkpointed_rounds_give_the_same_oracle(name):
    """CHECKPOINT_ROUNDS (used by the GPU-evaluated full-size comparisons) recomputes each depth round in the backward
    pass: same arithmetic, so outputs are identical and gradients agree to fp64 rounding"""
    if name not in golden_names():
        pytest.skip("no such golden case")
    meta, z, tree, graph = load_golden(name)
    names = OUT_NAMES[meta["kind"]]
    res = []
    for flag in (False, True):
        eo.CHECKPOINT_ROUNDS = flag
        try:
            sd, outs = _oracle_run(meta, tree, graph, torch.float64)
            sum((o * torch.from_numpy(z["cot_" + n])).sum() for n, o in zip(names, outs)).backward()
        finally:
            eo.CHECKPOINT_ROUNDS = False
        res.append(([o.detach() for o in outs], {k: p.grad for k, p in sd.items()}))
    for a, b in zip(res[0][0], res[1][0]):
        assert torch.equal(a, b)
    for k, g in res[0][1].items():
        if g is None:
            assert res[1][1][k] is None
        else:
            assert _rel(res[1][1][k].numpy(), g.numpy()) <= 1e-13, k


def test_oracle_fp32_close_to_fp64():
    meta, z, tree, graph = load_golden("hier_lstm_deep")
    _, outs = _oracle_run(meta, tree, graph, torch.float32)
    for nme, o in zip(OUT_NAMES["hier"], outs):
        assert _rel(o.detach().double().numpy(), z["out_" + nme]) <= 1e-5


@pytest.mark.skipif(not reference_available(), reason="/root/reference not present on this box")
@pytest.mark.parametrize("kind,rnn", [("hier", "LSTM"), ("hier", "GRU"), ("motif", "LSTM"), ("motif", "GRU")])
def test_oracle_matches_live_reference(kind, rnn):
    ref = load_reference()
    meta, z, tree, graph = load_golden("hier_lstm_small")
    nm, na = meta["vocab"]
    H = 20
    sd = {k: v.double() for k, v in eo.random_state(kind, rnn, H, H, 5, n_motif=nm, n_attach=na).items()}
    torch.set_default_dtype(torch.float64)
    try:
        cls = ref.encoder.HierMPNEncoder if kind == "hier" else ref.encoder.MotifEncoder
        model = cls(FakePairVocab(nm, na), IdAtomVocab(), rnn, H, H, 3, 4, 0.0)
        assert list(model.state_dict().keys()) == list(sd.keys())
        assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == \
            eo.reference_state_shapes(kind, rnn, H, H, n_motif=nm, n_attach=na)
        model.load_state_dict(sd, strict=True)
        outs = model(tree, graph) if kind == "hier" else model(tree)
    finally:
        torch.set_default_dtype(torch.float32)
    o = eo.hier_encoder(sd, rnn, 3, 4, tree, graph) if kind == "hier" else eo.motif_encoder(sd, rnn, 3, tree)
    for a, b in zip(o, outs):
        assert _rel(a.detach().numpy(), b.detach().numpy()) <= 1e-12


def test_unsupported_cell_type():
    meta, z, tree, graph = load_golden("motif_lstm_small")
    with pytest.raises(ValueError):
        _oracle_run(dict(meta, rnn="RNN"), tree, graph, torch.float64)
