"""Shared checks of the incremental encoders (IncMPNEncoder / IncHierMPNEncoder / IncEncoder, sparse_forward,
index_scatter), run on CPU against the C-ABI emulator (tests/test_inc.py) and on the B200 against the CUDA library
(tests/test_gpu_inc.py).  Expected values: tests/golden_inc/*.npz made by the unmodified reference
(oracle/make_golden_inc.py), and the oracle on seeded inputs for sizes without a fixture."""
import json
import os

import numpy as np
import torch

import ggpm_b200
from ggpm_b200 import encoder as genc
from oracle import inc_oracle as io
from oracle import make_golden_inc as mg
from oracle.ref_loader import FakePairVocab, IdAtomVocab

GOLDEN_INC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_inc")
TOL_OUT, TOL_GRAD = 1e-5, 1e-5       # as tests/_checks.py: fp32 path vs the fp64 reference


def inc_golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_INC) if f.endswith(".npz"))


def load(name):
    z = np.load(os.path.join(GOLDEN_INC, name + ".npz"))
    return json.loads(bytes(z["meta"]).decode()), z


def rel(a, b):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-30)


def build_model(case, device):
    nm, na = case["vocab"]
    cls = genc.IncHierMPNEncoder if case["kind"] == "hier" else genc.IncEncoder
    model = cls(FakePairVocab(nm, na), IdAtomVocab(), case["rnn"], case["hidden"], case["hidden"], case["depthT"],
                case["depthG"], 0.0)
    sd = mg.inc_state(case)
    assert list(model.state_dict().keys()) == list(sd.keys())
    model.load_state_dict(sd, strict=True)
    return model.to(device).eval()


def run_product(case, device, max_steps=None):
    hier = case["kind"] == "hier"
    _, tree_t, graph_t, sched, init = mg.build_inputs(case, device)
    model = build_model(case, device)
    init_p = init.float().to(device).requires_grad_(True)
    outs = io.run_decode(model, hier, tree_t, graph_t, sched, init_p, device=device, max_steps=max_steps)
    return model, outs, init_p, (tree_t, graph_t, sched)


def check_golden(name, device):
    case, z = load(name)
    model, outs, init_p, (tree_t, graph_t, sched) = run_product(case, device)
    assert mg.inputs_digest(tree_t, graph_t, sched) == case["digest"], "regenerated inputs differ from the fixture's"
    assert len(outs) == case["n_outputs"]
    for i, o in enumerate(outs):
        ref = z["out_%03d" % i]
        assert o.dtype == torch.float32 and tuple(o.shape) == ref.shape, i
        if ref.size:
            assert rel(o, ref) <= TOL_OUT, (i, rel(o, ref))
    io.scalar_loss(outs, io.cotangents(outs, case["weight_seed"] + 7)).backward()
    for k, p in model.named_parameters():
        ref = z["grad_" + k]
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        if np.abs(ref).max() == 0:
            assert float(g.abs().max()) == 0.0, k
        else:
            assert rel(g, ref) <= TOL_GRAD, (k, rel(g, ref))
    assert rel(init_p.grad, z["grad___init_vecs"]) <= TOL_GRAD


def check_against_oracle(case, device, max_steps=None, tol_out=TOL_OUT, tol_grad=TOL_GRAD):
    """sizes without a fixture: fp64 oracle on the same seeded inputs"""
    hier = case["kind"] == "hier"
    model, outs, init_p, _ = run_product(case, device, max_steps)
    _, tree_c, graph_c, sched, init = mg.build_inputs(case)
    sdo = {k: v.double().requires_grad_(True) for k, v in mg.inc_state(case).items()}
    orc = io.IncHierOracle(sdo, case["rnn"], case["depthT"], case["depthG"]) if hier else \
        io.IncMotifOracle(sdo, case["rnn"], case["depthT"])
    init_o = init.clone().requires_grad_(True)
    o_outs = io.run_decode(orc, hier, tree_c, graph_c, sched, init_o, max_steps=max_steps)
    assert len(outs) == len(o_outs)
    for i, (a, b) in enumerate(zip(outs, o_outs)):
        if b.numel():
            assert rel(a, b) <= tol_out, (i, rel(a, b))
    cots = io.cotangents(o_outs, 5)
    io.scalar_loss(o_outs, cots).backward()
    io.scalar_loss(outs, [c.to(dtype=torch.float32, device=device) for c in cots]).backward()
    for k, p in model.named_parameters():
        ref = sdo[k].grad if sdo[k].grad is not None else torch.zeros_like(sdo[k])
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        if float(ref.abs().max()) == 0:
            assert float(g.abs().max()) == 0.0, k
        else:
            assert rel(g, ref) <= tol_grad, (k, rel(g, ref))
    assert rel(init_p.grad, init_o.grad) <= tol_grad


def check_index_scatter(device):
    """ggpm/nnutils.py:124-128 semantics incl. gradients of both arguments"""
    from ggpm_b200 import nnutils
    gen = torch.Generator().manual_seed(3)
    table = torch.randn(11, 6, generator=gen)
    sub = torch.randn(4, 6, generator=gen)
    idx = torch.tensor([7, 2, 10, 0])
    t_p, s_p = table.clone().to(device).requires_grad_(True), sub.clone().to(device).requires_grad_(True)
    out = nnutils.index_scatter(s_p, t_p, idx.to(device))
    t_o, s_o = table.double().requires_grad_(True), sub.double().requires_grad_(True)
    ref = io.index_scatter(s_o, t_o, idx)
    assert rel(out, ref) == 0.0
    cot = torch.randn(11, 6, generator=gen)
    (out * cot.to(device)).sum().backward()
    (ref * cot.double()).sum().backward()
    assert rel(t_p.grad, t_o.grad) == 0.0 and rel(s_p.grad, s_o.grad) == 0.0
    assert torch.equal(t_p.detach().cpu(), table), "index_scatter must not modify its inputs"
    empty = nnutils.index_scatter(s_p[:0], t_p, idx[:0].to(device))
    assert rel(empty, table) == 0.0


def check_sparse_forward_cell(rnn_type, device, H=10, IN=7, depth=3):
    """LSTM / GRU.sparse_forward on a hand-made table: slot 0 entries in the middle of rows, a sub-message that is its
    own neighbour's neighbour, state rows beyond the message count; vs the oracle in fp64, with all gradients"""
    gen = torch.Generator().manual_seed(17)
    R, M, K = 9, 3, 4
    cls = ggpm_b200.LSTM if rnn_type == "LSTM" else ggpm_b200.GRU
    cell = cls(IN, H, depth)
    for p in cell.parameters():
        p.data = torch.randn(p.shape, generator=gen) * 0.3
    cell = cell.to(device)
    sd = {"rnn." + k: v.detach().cpu().double().requires_grad_(True) for k, v in cell.state_dict().items()}
    h0, c0 = torch.randn(R, H, generator=gen), torch.randn(R, H, generator=gen)     # row 0 deliberately non-zero
    x = torch.randn(M, IN, generator=gen)
    submess = torch.tensor([4, 1, 6])
    bgraph = torch.tensor([[1, 0, 8, 0], [0, 0, 0, 0], [4, 3, 0, 6]])
    hp, cp, xp = (t.clone().to(device).requires_grad_(True) for t in (h0, c0, x))
    ho, co, xo = (t.double().requires_grad_(True) for t in (h0, c0, x))
    if rnn_type == "LSTM":
        out = cell.sparse_forward((hp, cp), xp, submess.to(device), bgraph.to(device))
        ref = io.sparse_forward(sd, "rnn.", "LSTM", (ho, co), xo, submess, bgraph, depth)
    else:
        out = (cell.sparse_forward(hp, xp, submess.to(device), bgraph.to(device)),)
        ref = (io.sparse_forward(sd, "rnn.", "GRU", ho, xo, submess, bgraph, depth),)
    cots = [torch.randn(R, H, generator=gen) for _ in ref]
    for a, b in zip(out, ref):
        assert rel(a, b) <= TOL_OUT, rel(a, b)
    sum((a * c.to(device)).sum() for a, c in zip(out, cots)).backward()
    sum((b * c.double()).sum() for b, c in zip(ref, cots)).backward()
    pairs = [(hp, ho), (xp, xo)] + ([(cp, co)] if rnn_type == "LSTM" else [])
    for a, b in pairs:
        assert rel(a.grad, b.grad) <= TOL_GRAD, rel(a.grad, b.grad)
    for k, p in cell.named_parameters():
        assert rel(p.grad, sd["rnn." + k].grad) <= TOL_GRAD, (k, rel(p.grad, sd["rnn." + k].grad))
    assert torch.equal(hp.detach().cpu(), h0), "sparse_forward must not modify the incoming table"


def check_packed_weight_reuse(device):
    """sparse_forward shares one packing of the gate weights between calls while the parameters are unchanged: two
    calls inside one graph, a second forward + backward after the first graph was consumed, and an in-place update of a
    parameter (as an optimizer step does) must all give the oracle's values and gradients"""
    gen = torch.Generator().manual_seed(23)
    H, IN, R = 12, 6, 8
    cell = ggpm_b200.LSTM(IN, H, 2)
    for p in cell.parameters():
        p.data = torch.randn(p.shape, generator=gen) * 0.3
    cell = cell.to(device)
    h0, c0 = torch.randn(R, H, generator=gen), torch.randn(R, H, generator=gen)
    x1, x2 = torch.randn(2, IN, generator=gen), torch.randn(3, IN, generator=gen)
    s1, b1 = torch.tensor([3, 5]), torch.tensor([[1, 0, 2], [3, 0, 0]])
    s2, b2 = torch.tensor([1, 2, 6]), torch.tensor([[3, 5, 0], [0, 0, 0], [5, 1, 7]])

    def product():
        st = (h0.clone().to(device), c0.clone().to(device))
        st = cell.sparse_forward(st, x1.to(device), s1.to(device), b1.to(device))
        st = cell.sparse_forward(st, x2.to(device), s2.to(device), b2.to(device))
        cell.zero_grad(set_to_none=True)
        (st[0].sum() + 2 * st[1].sum()).backward()
        return st, {k: p.grad.detach().cpu().clone() for k, p in cell.named_parameters()}

    def oracle():
        sd = {"rnn." + k: v.detach().cpu().double().requires_grad_(True) for k, v in cell.state_dict().items()}
        st = (h0.double(), c0.double())
        st = io.sparse_forward(sd, "rnn.", "LSTM", st, x1.double(), s1, b1, 2)
        st = io.sparse_forward(sd, "rnn.", "LSTM", st, x2.double(), s2, b2, 2)
        (st[0].sum() + 2 * st[1].sum()).backward()
        return st, {k[4:]: v.grad for k, v in sd.items()}

    def same(a, b):
        (sa, ga), (sb, gb) = a, b
        assert rel(sa[0], sb[0]) <= TOL_OUT and rel(sa[1], sb[1]) <= TOL_OUT
        for k in gb:
            assert rel(ga[k], gb[k]) <= TOL_GRAD, (k, rel(ga[k], gb[k]))

    same(product(), oracle())
    from ggpm_b200 import rnn as grnn
    pack_a = grnn._STEP_PACKS[cell][1][0]
    same(product(), oracle())                               # second loop on unchanged weights reuses the packing
    assert grnn._STEP_PACKS[cell][1][0] is pack_a
    with torch.no_grad():
        cell.W_f[0].weight.add_(0.05)                       # what an optimizer step does
        cell.W[0].bias.mul_(-1.0)
    same(product(), oracle())
    assert grnn._STEP_PACKS[cell][1][0] is not pack_a
    import copy
    copy.deepcopy(cell)                                     # the cache must not ride along with the module
    with torch.no_grad():                                   # no graph in no_grad mode: separate packing
        out = cell.sparse_forward((h0.clone().to(device), c0.clone().to(device)), x1.to(device), s1.to(device), b1.to(device))
    assert not out[0].requires_grad


def check_readout_linear_cache_invalidation(device):
    """the padded read-out weights are shared between steps; an in-place parameter update must be picked up"""
    case, _ = load("inc_motif_lstm")
    _, tree_t, graph_t, sched, init = mg.build_inputs(case, device)
    model = build_model(case, device)
    init_d = init.float().to(device)
    with torch.no_grad():
        first = io.run_decode(model, False, tree_t, graph_t, sched, init_d, device=device, max_steps=3)
        model.tree_encoder.W_o[0].weight.mul_(1.5)
        model.tree_encoder.W_o[0].bias.add_(0.25)
        second = io.run_decode(model, False, tree_t, graph_t, sched, init_d, device=device, max_steps=3)
    _, tree_c, graph_c, _, _ = mg.build_inputs(case)
    sd = {k: v.detach().cpu().double() for k, v in model.state_dict().items()}
    want = io.run_decode(io.IncMotifOracle(sd, case["rnn"], case["depthT"]), False, tree_c, graph_c, sched, init, max_steps=3)
    assert rel(first[0], want[0]) > 1e-3
    for a, b in zip(second, want):
        if b.numel():
            assert rel(a, b) <= TOL_OUT, rel(a, b)
