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
# fp32 path vs the fp64 reference run: the north-star bar, 1e-5 of max|ref|, for outputs AND gradients.
# The one documented exception (`grad_tolerance`): a parameter gradient whose value the REFERENCE's own fp32 run
# cannot reproduce to 1e-5 (a ReLU pre-activation within rounding of zero flips that unit's whole downstream
# gradient) is held to twice the measured reference-fp32-vs-fp64 gap of that tensor instead.
TOL_OUT, TOL_GRAD = 1e-5, 1e-5

# measured worst case per tensor of every parity comparison of the session: {test: {tensor: {"err", "tol", ...}}};
# tests/conftest.py writes it to gpurun_out/parity_report.json at session end on a GPU box
REPORT = {}


def record(test, tensor, err, tol, **extra):
    REPORT.setdefault(test, {})[tensor] = dict(err=float(err), tol=float(tol), **extra)


def grad_tolerance(gap32, bar=None):
    """gap32: rel(reference fp32 gradient, reference fp64 gradient) of one tensor, or None when not measured;
    bar: the case's stated gradient bar when it is not the north star's TOL_GRAD (documented where it is passed)"""
    bar = TOL_GRAD if bar is None else bar
    return bar if gap32 is None else max(bar, 2.0 * gap32)


def compare_with_oracle(test, names, outs, ref_outs, grads, ref_grads, ref_grads32=None, ref_outs32=None, raise_on_bad=True,
                        skip_outputs=False, grad_bar=None):
    """outputs at TOL_OUT, every parameter gradient at TOL_GRAD — each relaxed to twice the reference's OWN fp32-vs-fp64
    gap on that tensor where that gap was measured and is larger (grad_tolerance); everything recorded.  Returns the
    list of violations when raise_on_bad is False."""
    bad = []
    for i, (n, a, b) in enumerate(zip(names, outs, ref_outs)):
        if skip_outputs:
            break
        e = rel(a, b)
        gap = None if ref_outs32 is None else rel(ref_outs32[i], b)
        tol = grad_tolerance(gap)
        record(test, "out:" + n, e, tol, ref_fp32_gap=gap)
        if e > tol:
            bad.append((n, e, tol))
    for k, g in ref_grads.items():
        if g is None or float(g.abs().max()) == 0.0:
            assert k not in grads or grads[k] is None or float(grads[k].abs().max()) == 0.0, k
            continue
        assert grads.get(k) is not None, "no gradient for " + k
        gap = None if ref_grads32 is None or ref_grads32.get(k) is None else rel(ref_grads32[k], g)
        tol = grad_tolerance(gap, grad_bar)
        e = rel(grads[k], g)
        record(test, "grad:" + k, e, tol, ref_fp32_gap=gap)
        if e > tol:
            bad.append((k, e, tol))
    if raise_on_bad:
        assert not bad, bad
    return bad


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
    tag = "golden:%s%s:engine%s" % (name, ":packed" if packed else "", _engine_tag())
    for nme, o in zip(names, outs):
        assert o.dtype == torch.float32 and o.is_contiguous()
        assert tuple(o.shape) == z["out_" + nme].shape
        record(tag, "out:" + nme, rel(o, z["out_" + nme]), TOL_OUT)
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
            record(tag, "grad:" + k, rel(p.grad, ref), TOL_GRAD)
            assert rel(p.grad, ref) <= TOL_GRAD, (k, rel(p.grad, ref))


def _engine_tag():
    from ggpm_b200 import _lib
    try:
        return str(int(_lib.backend().ggpm_get_gemm_engine()))
    except Exception:
        return "?"


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
    """train mode with the real nn.Dropout modules: two calls differ, gradients are finite, and at the first dropout
    site of the path (the atom read-out W_o, ggpm/encoder.py:15-19,33-34 — nothing upstream of it is random) every
    element is either dropped or the eval-mode value / (1 - p), with a drop rate near p"""
    meta, z, tree, graph = load_golden("hier_lstm_small")
    tree, graph = to_dev(tree, device), to_dev(graph, device)
    p = 0.5
    model = build_model(meta, device, dropout=p).train()
    torch.manual_seed(0)
    outs = model(tree, graph)
    a = outs[0]
    b = model(tree, graph)[0]
    assert not torch.equal(a, b)
    a.sum().backward()
    assert all(torch.isfinite(q.grad).all() for q in model.parameters() if q.grad is not None)
    hatom_train = outs[3].detach()
    with torch.no_grad():
        hatom_eval = model.eval()(tree, graph)[3]
    live = hatom_eval > 0                                   # ReLU zeros say nothing about dropout
    kept = hatom_train != 0
    scaled = hatom_eval / (1.0 - p)
    assert torch.allclose(hatom_train[kept], scaled[kept], rtol=1e-5, atol=1e-6)
    assert not bool((kept & ~live).any())
    rate = 1.0 - float(kept[live].float().mean())
    assert abs(rate - p) < 0.05, rate


class FixedDrop(torch.nn.Module):
    """stands in for an nn.Dropout of the model under test: multiplies by a fixed multiplier (mask / (1 - p)); the
    product carries node tables at a padded width, extra columns get multiplier 0"""

    p = 0.5         # any non-zero rate: callers that ask a dropout module for its multiplier only skip p == 0

    def __init__(self, mult):
        super().__init__()
        self.mult = mult

    def forward(self, x):
        m = self.mult.to(x.device, x.dtype)
        if m.shape[1] < x.shape[1]:
            m = torch.nn.functional.pad(m, (0, x.shape[1] - m.shape[1]))
        return x * m


def install_fixed_dropout(model, kind, tree, graph, hidden, embed, p, seed):
    """replaces every dropout module of a ggpm_b200 encoder by FixedDrop with seeded masks; returns the oracle's
    `drop` dict (same multipliers, fp64)"""
    gen = torch.Generator().manual_seed(seed)
    n_t = tree[0].shape[0]
    sites = {"E_i": (n_t, embed), "E_c": (n_t, embed), "tree_encoder.W_o": (n_t, hidden)}
    if kind == "hier":
        sites.update({"W_i": (n_t, hidden), "W_c": (n_t, hidden), "inter_encoder.W_o": (n_t, hidden),
                      "graph_encoder.W_o": (graph[0].shape[0], hidden)})
    drop = {k: (torch.rand(shape, generator=gen) >= p).double() / (1.0 - p) for k, shape in sites.items()}
    model.E_i = torch.nn.Sequential(model.E_i[0], FixedDrop(drop["E_i"]))
    model.E_c = torch.nn.Sequential(model.E_c[0], FixedDrop(drop["E_c"]))
    model.tree_encoder.W_o[2] = FixedDrop(drop["tree_encoder.W_o"])
    if kind == "hier":
        model.W_i[2] = FixedDrop(drop["W_i"])
        model.W_c[2] = FixedDrop(drop["W_c"])
        model.inter_encoder.W_o[2] = FixedDrop(drop["inter_encoder.W_o"])
        model.graph_encoder.W_o[2] = FixedDrop(drop["graph_encoder.W_o"])
    return drop


def check_fixed_mask_dropout_parity(name, device, p=0.1):
    """train-mode parity (the benchmarked cfg2 / cfg5 run dropout 0.1 in train mode): the same fixed dropout
    multipliers at every dropout site on both sides — the product with its nn.Dropout modules replaced by FixedDrop,
    the fp64 oracle fed the same masks where the reference's nn.Dropout modules sit — outputs and all gradients"""
    meta, z, tree, graph = load_golden(name)
    kind, H = meta["kind"], meta["hidden"]
    model = build_model(meta, device, dropout=p).train()
    drop = install_fixed_dropout(model, kind, tree, graph, H, H, p, seed=17)
    outs = run(model, meta, to_dev(tree, device), to_dev(graph, device))
    names = OUT_NAMES[kind]
    cots = [torch.from_numpy(z["cot_" + n]) for n in names]
    sum((o * c.float().to(device)).sum() for o, c in zip(outs, cots)).backward()
    nm, na = meta["vocab"]
    sd = eo.random_state(kind, meta["rnn"], H, H, meta["weight_seed"], n_motif=nm, n_attach=na)
    sdo = {k: v.double().requires_grad_(True) for k, v in sd.items()}
    if kind == "hier":
        ref = eo.hier_encoder(sdo, meta["rnn"], meta["depthT"], meta["depthG"], tree, graph, drop=drop)
    else:
        ref = eo.motif_encoder(sdo, meta["rnn"], meta["depthT"], tree, drop=drop)
    sum((o * c.double()).sum() for o, c in zip(ref, cots)).backward()
    # the masks must matter: the dropout-free golden outputs differ
    assert rel(ref[0], z["out_" + names[0]]) > 1e-3
    compare_with_oracle("fixed_mask_dropout:%s:p%.1f" % (name, p), names, outs, ref,
                        {k: q.grad for k, q in model.named_parameters()}, {k: v.grad for k, v in sdo.items()})


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


def check_malformed_batches_raise(device):
    """a bad label / position / source node / neighbour id must raise IndexError (as the reference's index_select
    does) instead of reaching a kernel: on the make_cuda path through the device-side IndexGuard, on the ingest path
    on the host when the batch is packed"""
    from ggpm_b200 import ingest
    meta, z, tree, graph = load_golden("hier_lstm_small")
    model = build_model(meta, device)

    def corrupt(which, col, value):
        t, g = [x.clone() if isinstance(x, torch.Tensor) else x for x in tree], [x.clone() if isinstance(x, torch.Tensor) else x for x in graph]
        tgt = (t if which[0] == "t" else g)[{"fmess": 1, "agraph": 2, "bgraph": 3, "cgraph": 4}[which[1:]]]
        if col is None:
            tgt[1, 0] = value
        else:
            tgt[1, col] = value
        return t, g

    cases = [("tfmess", 2, 20), ("tfmess", 0, tree[0].shape[0]), ("gfmess", 2, 4), ("gfmess", 3, 20), ("gfmess", 0, graph[0].shape[0]),
             ("tbgraph", None, tree[1].shape[0]), ("gagraph", None, graph[1].shape[0]), ("tcgraph", None, graph[0].shape[0]),
             ("gbgraph", None, -1)]
    for which, col, value in cases:
        t, g = corrupt(which, col, value)
        with pytest.raises(IndexError):
            model(to_dev(t, device), to_dev(g, device))
        with pytest.raises(IndexError):
            ingest.PackedBatch.from_tensors(t, g)
    # the untouched batch still runs
    model(to_dev(tree, device), to_dev(graph, device))


def check_lstm_cell_state_gradient(device):
    """LSTM.forward returns (h, c) as the reference does (ggpm/rnn.py:108): a loss on c must reach the parameters and
    the inputs, and an in-place edit of the returned tensors must not corrupt the backward pass"""
    meta, z, tree, graph = load_golden("motif_lstm_small")
    H, IN, depth = 12, 9, 4
    torch.manual_seed(4)
    rnn = ggpm_b200.LSTM(IN, H, depth)
    for p in rnn.parameters():
        torch.nn.init.normal_(p, std=0.3)
    rnn = rnn.to(device)
    bgraph = tree[3]
    E1 = tree[1].shape[0]
    x = torch.randn(E1, IN)
    ca, cb = torch.randn(E1, H), torch.randn(E1, H)
    xd = x.clone().to(device).requires_grad_(True)
    h, c = rnn(xd, bgraph.to(device))
    loss = (h * ca.to(device)).sum() + (c * cb.to(device)).sum()
    with torch.no_grad():          # the returned states are the caller's: scribbling on them must be harmless
        h.zero_()
        c.zero_()
    loss.backward()
    sd = {"rnn." + k: v.detach().cpu().double().requires_grad_(True) for k, v in rnn.state_dict().items()}
    xo = x.clone().double().requires_grad_(True)
    ho, co = eo.lstm_messages(sd, "rnn.", xo, bgraph, depth)
    ((ho * ca.double()).sum() + (co * cb.double()).sum()).backward()
    assert rel(xd.grad, xo.grad) <= TOL_GRAD, rel(xd.grad, xo.grad)
    for k, p in rnn.named_parameters():
        assert rel(p.grad, sd["rnn." + k].grad) <= TOL_GRAD, (k, rel(p.grad, sd["rnn." + k].grad))
    # c alone
    rnn.zero_grad(set_to_none=True)
    rnn(x.to(device), bgraph.to(device))[1].sum().backward()
    for v in sd.values():
        v.grad = None
    eo.lstm_messages(sd, "rnn.", x.double(), bgraph, depth)[1].sum().backward()
    for k, p in rnn.named_parameters():
        assert rel(p.grad, sd["rnn." + k].grad) <= TOL_GRAD, (k, rel(p.grad, sd["rnn." + k].grad))
