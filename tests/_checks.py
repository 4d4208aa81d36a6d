"""Shared encoder-level checks, run on CPU against the C-ABI emulator (tests/test_host_emulated.py) and on
the B200 against the real CUDA library (tests/test_gpu_encoder.py).  Expected values: golden vectors made
by the unmodified reference (oracle/make_golden.py) and the oracle on the same seeded inputs."""
import numpy as np
import pytest
import torch

from conftest import load_golden
import ggpm_b200
from ggpm_b200 import nnutils
from oracle import encoder_oracle as eo
from oracle.ref_loader import FakePairVocab, IdAtomVocab

OUT_NAMES = {"hier": ("hroot", "hnode", "hinter", "hatom"), "motif": ("hroot", "hnode")}
# fp32 path vs fp64 reference: the north-star bar (1e-5 relative) for outputs; gradients get 2e-5 because
# the reference's own fp32 run differs from its fp64 run by up to ~1e-6 per layer
TOL_OUT, TOL_GRAD = 1e-5, 2e-5


def build_model(meta, device, dropout=0.0):
    nm, na = meta["vocab"]
    cls = ggpm_b200.HierMPNEncoder if meta["kind"] == "hier" else ggpm_b200.MotifEncoder
    model = cls(FakePairVocab(nm, na), IdAtomVocab(), meta["rnn"], meta["hidden"], meta["hidden"], meta["depthT"],
                meta["depthG"], dropout)
    sd = eo.random_state(meta["kind"], meta["rnn"], meta["hidden"], meta["hidden"], meta["weight_seed"],
                         n_motif=nm, n_attach=na)
    assert list(model.state_dict().keys()) == list(sd.keys())
    model.load_state_dict(sd, strict=True)
    return model.to(device)


def to_dev(tensors, device):
    return [t.to(device) if isinstance(t, torch.Tensor) else t for t in tensors]


def rel(a, b):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-30)


def run(model, meta, tree, graph):
    return model(tree, graph) if meta["kind"] == "hier" else model(tree)


def check_forward_backward_golden(name, device, packed=False):
    meta, z, tree, graph = load_golden(name)
    if packed:      # through ingest.PackedBatch: one buffer, int32 views, index structures prepared on the host
        from ggpm_b200 import ingest
        tree, graph = ingest.PackedBatch.from_tensors(tree, graph if meta["kind"] == "hier" else None).to_device(device)
    else:
        tree, graph = to_dev(tree, device), to_dev(graph, device)
    model = build_model(meta, device)
    outs = run(model, meta, tree, graph)
    names = OUT_NAMES[meta["kind"]]
    for nme, o in zip(names, outs):
        assert o.dtype == torch.float32 and o.is_contiguous()
        assert tuple(o.shape) == z["out_" + nme].shape
        assert rel(o, z["out_" + nme]) <= TOL_OUT, (nme, rel(o, z["out_" + nme]))
    for o in outs[1:]:
        assert float(o[0].detach().abs().max()) == 0.0
    sum((o * torch.from_numpy(z["cot_" + n]).float().to(device)).sum() for n, o in zip(names, outs)).backward()
    for k, p in model.named_parameters():
        ref = z["gall_" + k]
        if np.abs(ref).max() == 0:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
        else:
            assert p.grad is not None, k
            assert rel(p.grad, ref) <= TOL_GRAD, (k, rel(p.grad, ref))


def check_root_only_loss(device):
    meta, z, tree, graph = load_golden("hier_lstm_small")
    model = build_model(meta, device)
    model(to_dev(tree, device), to_dev(graph, device))[0].sum().backward()
    for k, p in model.named_parameters():
        if "groot_" + k in z.files:
            assert rel(p.grad, z["groot_" + k]) <= TOL_GRAD, k
        else:
            assert k.startswith("tree_encoder.W_o") and p.grad is None


def check_cfg1(device):
    """BASELINE config #1 stand-in: H=250, depth 20/20, 20 molecules; outputs + strided hroot.sum() grads"""
    meta, z, tree, graph = load_golden("hier_lstm_cfg1")
    model = build_model(meta, device)
    outs = model(to_dev(tree, device), to_dev(graph, device))
    for nme, o in zip(OUT_NAMES["hier"], outs):
        assert rel(o, z["out_" + nme]) <= TOL_OUT, (nme, rel(o, z["out_" + nme]))
    outs[0].sum().backward()
    stride = meta["grad_stride"]
    for k, p in model.named_parameters():
        if "groot_" + k not in z.files:
            assert p.grad is None
            continue
        g = p.grad.reshape(-1)[::stride]
        scale = float(z["gnorm_" + k]) / max(p.grad.numel() ** 0.5, 1.0)      # rms of the full gradient
        err = float(np.abs(g.cpu().numpy().astype(np.float64) - z["groot_" + k]).max())
        assert err <= TOL_GRAD * max(float(np.abs(z["groot_" + k]).max()), scale), (k, err)
        assert abs(float(p.grad.double().norm()) - float(z["gnorm_" + k])) <= TOL_GRAD * float(z["gnorm_" + k]), k


def check_no_grad_eval(device):
    meta, z, tree, graph = load_golden("hier_gru_small")
    model = build_model(meta, device, dropout=0.3).eval()
    with torch.no_grad():
        outs = model(to_dev(tree, device), to_dev(graph, device))
    for nme, o in zip(OUT_NAMES["hier"], outs):
        assert rel(o, z["out_" + nme]) <= TOL_OUT


def check_train_dropout(device):
    meta, z, tree, graph = load_golden("hier_lstm_small")
    tree, graph = to_dev(tree, device), to_dev(graph, device)
    model = build_model(meta, device, dropout=0.5).train()
    torch.manual_seed(0)
    a = model(tree, graph)[0]
    b = model(tree, graph)[0]
    assert not torch.equal(a, b)
    a.sum().backward()
    assert all(torch.isfinite(p.grad).all() for p in model.parameters() if p.grad is not None)


def check_inputs_not_mutated(device):
    meta, z, tree, graph = load_golden("motif_lstm_small")
    tree = to_dev(tree, device)
    model = build_model(meta, device)
    before = [t.clone() for t in tree[:-1]]
    model(tree)
    for a, b in zip(before, tree[:-1]):
        assert torch.equal(a, b)


def check_mpn_reference_signature(rnn, device):
    """MPNEncoder.forward(fnode, fmess, agraph, bgraph) with DENSE message inputs (ggpm/encoder.py:28-38)
    and the rnn modules' own forward(fmess, bgraph) (ggpm/rnn.py:41-50, 96-108)."""
    meta, z, tree, graph = load_golden("motif_lstm_small")
    H, IN, Fd, depth = 12, 9, 7, 4
    torch.manual_seed(3)
    mpn = ggpm_b200.MPNEncoder(rnn, IN, Fd, H, depth, 0.0)
    for p in mpn.parameters():
        torch.nn.init.normal_(p, std=0.3)
    mpn = mpn.to(device)
    fnode = torch.randn(tree[0].shape[0], Fd)
    fmess = torch.randn(tree[1].shape[0], IN)
    agraph, bgraph = tree[2], tree[3]
    node, h = mpn(fnode.to(device), fmess.to(device), agraph.to(device), bgraph.to(device))
    sd = {k: v.detach().cpu().double() for k, v in mpn.state_dict().items()}
    o_node, o_h = eo.mpn_layer(sd, "", rnn, fnode.double(), fmess.double(), agraph, bgraph, depth)
    assert rel(node, o_node) <= TOL_OUT
    assert rel(h, o_h) <= TOL_OUT
    out = mpn.rnn(fmess.to(device), bgraph.to(device))
    hh = out[0] if rnn == "LSTM" else out
    assert rel(hh, o_h) <= TOL_OUT
    if rnn == "LSTM":
        _, o_c = eo.lstm_messages(sd, "rnn.", fmess.double(), bgraph, depth)
        assert rel(out[1], o_c) <= TOL_OUT
    with pytest.raises(ValueError):
        ggpm_b200.MPNEncoder("RNN", IN, Fd, H, depth, 0.0)


def check_index_select_nd(device):
    src = torch.randn(7, 8).to(device)
    idx = torch.tensor([[1, 2, 0], [0, 0, 0], [6, 6, 3]]).to(device)
    got = nnutils.index_select_ND(src, 0, idx)
    assert torch.allclose(got, src[idx.reshape(-1)].reshape(3, 3, 8))
    assert torch.allclose(nnutils.index_select_sum(src, idx), (src[idx] * (idx != 0).unsqueeze(-1)).sum(1))
