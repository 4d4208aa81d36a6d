"""CPU: the incremental encoders.  (1) the oracle restatement against the golden vectors made by
the unmodified reference, and against the live reference when /root/reference exists; (2) the product's host side
(sub-tensor selection, weight slicing, autograd wiring, pointer / stride arithmetic of every C-ABI call) driven
through the C-ABI emulator and compared with the same golden vectors."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _emu_lib                                                        # noqa: E402
import _inc_checks as IC                                               # noqa: E402
from ggpm_b200 import _lib, nnutils                                    # noqa: E402
from oracle import inc_oracle as io                                    # noqa: E402
from oracle import make_golden_inc as mg                               # noqa: E402
from oracle.ref_loader import FakePairVocab, IdAtomVocab, load_reference, reference_available  # noqa: E402

NAMES = IC.inc_golden_names()


@pytest.fixture()
def emu():
    lib = _emu_lib.EmuLib()
    _lib._install_test_backend(lib)
    nnutils.set_default_device("cpu")
    yield lib
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)


def test_fixtures_present():
    assert set(NAMES) >= {"inc_hier_lstm", "inc_hier_gru", "inc_motif_lstm", "inc_motif_gru", "inc_hier_lstm_diter"}


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_golden(name):
    case, z = IC.load(name)
    hier = case["kind"] == "hier"
    _, tree_t, graph_t, sched, init = mg.build_inputs(case)
    assert mg.inputs_digest(tree_t, graph_t, sched) == case["digest"]
    sdo = {k: v.double().requires_grad_(True) for k, v in mg.inc_state(case).items()}
    orc = io.IncHierOracle(sdo, case["rnn"], case["depthT"], case["depthG"]) if hier else \
        io.IncMotifOracle(sdo, case["rnn"], case["depthT"])
    init_o = init.clone().requires_grad_(True)
    outs = io.run_decode(orc, hier, tree_t, graph_t, sched, init_o)
    assert len(outs) == case["n_outputs"]
    for i, o in enumerate(outs):
        if o.numel():
            assert IC.rel(o, z["out_%03d" % i]) <= 1e-12, i
    io.scalar_loss(outs, io.cotangents(outs, case["weight_seed"] + 7)).backward()
    for k, v in sdo.items():
        ref = z["grad_" + k]
        g = v.grad if v.grad is not None else torch.zeros_like(v)
        if np.abs(ref).max() > 0:
            assert IC.rel(g, ref) <= 1e-12, k
    assert IC.rel(init_o.grad, z["grad___init_vecs"]) <= 1e-12


@pytest.mark.skipif(not reference_available(), reason="reference checkout not present (GPU box)")
def test_oracle_matches_live_reference_on_another_seed():
    case = dict(name="live", kind="hier", rnn="LSTM", hidden=16, depthT=2, depthG=2, batch=3, data_seed=77,
                weight_seed=78, vocab=(30, 30), mean_motifs=4.0)
    ref = load_reference()
    _, tree_t, graph_t, sched, init = mg.build_inputs(case)
    sd64 = mg.inc_state(case, torch.float64)
    torch.set_default_dtype(torch.float64)
    try:
        model = ref.encoder.IncHierMPNEncoder(FakePairVocab(30, 30), IdAtomVocab(), "LSTM", 16, 16, 2, 2, 0.0)
        model.load_state_dict(sd64, strict=True)
        outs = io.run_decode(model.double().eval(), True, tree_t, graph_t, sched, init)
    finally:
        torch.set_default_dtype(torch.float32)
    o_outs = io.run_decode(io.IncHierOracle(sd64, "LSTM", 2, 2), True, tree_t, graph_t, sched, init)
    assert max(IC.rel(a, b) for a, b in zip(o_outs, outs) if b.numel()) <= 1e-12


def test_schedule_covers_every_tree_message_once_down_once_up():
    case, _ = IC.load("inc_hier_lstm")
    _, tree_t, graph_t, sched, _ = mg.build_inputs(case)
    mess = [m for st in sched["steps"] for m in st["mess"]]
    assert sorted(mess) == list(range(1, tree_t[1].shape[0]))          # every directed tree message exactly once
    assert all(len(st["nodes"]) >= len(st["mess"]) for st in sched["steps"])
    atoms = set(sched["first_atoms"]) | {a for st in sched["steps"] for a in st["next_atoms"]}
    assert atoms == set(range(1, graph_t[0].shape[0]))                 # every atom is eventually switched on


@pytest.mark.parametrize("name", NAMES)
def test_product_host_side_matches_reference_golden(emu, name):
    IC.check_golden(name, "cpu")
    assert "rows_gather" in emu.calls and "rows_scatter" in emu.calls
    assert ("sparse_lstm_cell_bwd" in emu.calls) or ("sparse_gru_gate_bwd" in emu.calls)


def test_index_scatter(emu):
    IC.check_index_scatter("cpu")


@pytest.mark.parametrize("rnn", ["LSTM", "GRU"])
def test_sparse_forward_cell(emu, rnn):
    IC.check_sparse_forward_cell(rnn, "cpu")


def test_state_dict_interchanges_with_reference_keys(emu):
    case, _ = IC.load("inc_hier_gru")
    model = IC.build_model(case, "cpu")
    assert not any(k.startswith("W_root") for k in model.state_dict())
    assert type(model.tree_encoder).__name__ == "IncMPNEncoder" and hasattr(model, "embed_graph")


def test_sparse_forward_rejects_mismatched_arguments(emu):
    import ggpm_b200
    cell = ggpm_b200.LSTM(5, 8, 2)
    h = (torch.zeros(6, 8), torch.zeros(6, 8))
    with pytest.raises(ValueError):
        cell.sparse_forward(h, torch.zeros(2, 4), torch.tensor([1, 2]), torch.zeros(2, 3, dtype=torch.long))
    with pytest.raises(ValueError):
        cell.sparse_forward(h, torch.zeros(2, 5), torch.tensor([1, 2, 3]), torch.zeros(2, 3, dtype=torch.long))


def test_packed_weights_are_reused_until_parameters_change(emu):
    IC.check_packed_weight_reuse("cpu")


@pytest.mark.skipif(not reference_available(), reason="reference checkout not present (GPU box)")
@pytest.mark.parametrize("kind,rnn", [("hier", "LSTM"), ("hier", "GRU"), ("motif", "LSTM")])
def test_state_dict_interchanges_with_live_reference(emu, kind, rnn):
    """same keys, shapes and registration order as the reference's Inc classes; strict loads both ways"""
    from ggpm_b200 import encoder as genc
    ref = load_reference()
    args = (FakePairVocab(30, 40), IdAtomVocab(), rnn, 16, 16, 2, 3, 0.0)
    r = (ref.encoder.IncHierMPNEncoder if kind == "hier" else ref.encoder.IncEncoder)(*args)
    m = (genc.IncHierMPNEncoder if kind == "hier" else genc.IncEncoder)(*args)
    rs, ms = r.state_dict(), m.state_dict()
    assert list(rs.keys()) == list(ms.keys())
    assert [tuple(v.shape) for v in rs.values()] == [tuple(v.shape) for v in ms.values()]
    assert [n for n, _ in r.named_parameters()] == [n for n, _ in m.named_parameters()]
    m.load_state_dict(rs, strict=True)
    r.load_state_dict(ms, strict=True)
    assert not hasattr(m, "W_root") and not hasattr(r, "W_root")
    for name in ("get_sub_tensor", "embed_sub_tree", "forward"):
        assert hasattr(m, name)


def test_tie_embedding_hands_over_the_decoder_side_modules(emu):
    """encoder.tie_embedding(decoder.hmpn) (ggpm/encoder.py:92-94): the encoder adopts the incremental encoder's E_c /
    E_i modules and one-hot tables, so both sides train the same embedding"""
    import ggpm_b200
    from ggpm_b200 import encoder as genc
    args = (FakePairVocab(30, 40), IdAtomVocab(), "LSTM", 16, 16, 2, 3, 0.0)
    hmpn = genc.IncHierMPNEncoder(*args)
    enc = ggpm_b200.HierMPNEncoder(*args)
    enc.tie_embedding(hmpn)
    assert enc.E_c is hmpn.E_c and enc.E_i is hmpn.E_i and enc.E_a is hmpn.E_a and enc.E_b is hmpn.E_b


def test_dropout_in_train_mode_reaches_the_incremental_readout(emu):
    case, _ = IC.load("inc_motif_lstm")
    from ggpm_b200 import encoder as genc
    nm, na = case["vocab"]
    model = genc.IncEncoder(FakePairVocab(nm, na), IdAtomVocab(), "LSTM", case["hidden"], case["hidden"], case["depthT"], 0, 0.5)
    model.load_state_dict(mg.inc_state(case), strict=True)
    _, tree_t, graph_t, sched, init = mg.build_inputs(case)
    torch.manual_seed(0)
    a = io.run_decode(model.train(), False, tree_t, graph_t, sched, init.float(), max_steps=2)
    b = io.run_decode(model.eval(), False, tree_t, graph_t, sched, init.float(), max_steps=2)
    assert not torch.allclose(a[0], b[0])            # node read-out goes through W_o's Dropout (ggpm/encoder.py:12-16)
    assert float((a[0] == 0).float().mean()) > 0.2


def test_readout_weights_are_repadded_after_a_parameter_update(emu):
    IC.check_readout_linear_cache_invalidation("cpu")
