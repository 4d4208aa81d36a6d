"""B200: the CUDA path against the fp64 oracle (itself pinned to the live reference at 1e-12, tests/test_oracle_golden.py)
on the configurations the performance claims stand on — the comparisons VERDICT round 1 found missing:

  * hidden 500, depth 20/20 (BASELINE config #3's width: Hp = 512, K = 512 tensor-core tiles);
  * hidden 250 with a batch whose atom layer crosses 32 768 message rows NATURALLY, so the forward takes the
    `lstm_sgather` + plane-fed GEMM path every cfg3 / cfg4 atom layer takes (no environment override);
  * MotifEncoder + the latent head at hidden 250, latent 24 (BASELINE config #5, the PropertyVAE encoder side);
  * the graph GRU end to end at hidden 250, depth 20.

Bar: outputs and every parameter gradient within 1e-5 of max|ref| (north star).  Two documented, MEASURED exceptions,
both recorded in gpurun_out/parity_report.json:

  * where the REFERENCE's own fp32 evaluation misses its fp64 evaluation by more than 5e-6 on a tensor, that tensor is
    held to twice the measured reference gap (tests/_checks.py `grad_tolerance`; e.g. the GRU, whose state is an
    unsquashed neighbour sum that grows with depth, ggpm/rnn.py:38);
  * ReLU kinks.  At these sizes (10^5-10^6 ReLU units) a few pre-activations sit within fp32 rounding of zero, and
    whichever side an evaluation lands on switches that unit's whole downstream gradient — the reference's own fp32 run
    has the same property.  When the plain comparison fails, the test (1) reads the activation pattern the product took
    (forward hooks on its five ReLU sites), (2) requires that every unit whose pattern differs from the fp64 oracle's
    has an fp64 pre-activation below 1e-5 of the site's scale, and that there are at most max(16, units / 250 000)
    of them, and (3) re-runs
    the fp64 oracle on the product's pattern (`relu=` of oracle/encoder_oracle.py: gradient of a prescribed linear
    branch) and holds every gradient to 1e-5 against THAT (relaxed, as above, to twice the gap of the reference's fp32
    arithmetic evaluated on the same branch where that gap exceeds 5e-6).  A wrong kernel fails (2) or (3); a kink
    passes both.
"""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _checks as C                                                                     # noqa: E402
import ggpm_b200                                                                       # noqa: E402
from ggpm_b200 import ingest, latent, synth, tensorize                                  # noqa: E402
from oracle import encoder_oracle as eo                                                # noqa: E402
from oracle import latent_oracle as lo                                                 # noqa: E402

pytestmark = pytest.mark.gpu


def _as_long(ts):
    return [torch.from_numpy(np.asarray(t)).long() if not isinstance(t, list) else t for t in ts]


def _oracle(kind, rnn, depthT, depthG, sd, tree_np, graph_np, dtype, loss, relu=None, device="cpu"):
    """-> (outputs, {param: grad}, {site: pre-activation}) of the oracle at `dtype` for loss(outs).  device="cuda"
    evaluates the SAME torch restatement with its tensors on the GPU (fp64 torch ops, depth rounds recomputed in the
    backward pass) — the only way the sweep-size batches finish in seconds; results come back on the host."""
    torch.set_num_threads(os.cpu_count() or 1)
    on = lambda t: t.to(device) if isinstance(t, torch.Tensor) else t
    leaves = {k: v.detach().clone().to(dtype).to(device).requires_grad_(True) for k, v in sd.items()}
    tree, graph = [on(t) for t in _as_long(tree_np)], [on(t) for t in _as_long(graph_np)]
    if relu is not None:
        relu = {k: v.to(device) for k, v in relu.items()}
    trace = {}
    eo.CHECKPOINT_ROUNDS = device != "cpu"
    try:
        if kind == "hier":
            outs = eo.hier_encoder(leaves, rnn, depthT, depthG, tree, graph, relu=relu, trace=trace)
        else:
            outs = eo.motif_encoder(leaves, rnn, depthT, tree, relu=relu, trace=trace)
        loss(outs, leaves).backward()
    finally:
        eo.CHECKPOINT_ROUNDS = False
    res = ([o.detach().cpu() for o in outs], {k: None if v.grad is None else v.grad.cpu() for k, v in leaves.items()},
           {k: v.cpu() for k, v in trace.items()})
    del outs, leaves, trace
    if device != "cpu":
        torch.cuda.empty_cache()
    return res


def _relu_sites(model, kind):
    """site name (oracle/encoder_oracle.py) -> the nn.Dropout module that receives that site's post-ReLU activations"""
    sites = {"tree_encoder.W_o": model.tree_encoder.W_o[2]}
    if kind == "hier":
        sites.update({"W_i": model.W_i[2], "W_c": model.W_c[2], "inter_encoder.W_o": model.inter_encoder.W_o[2],
                      "graph_encoder.W_o": model.graph_encoder.W_o[2]})
    return sites


def _product(kind, rnn, H, depthT, depthG, sd, mols, loss, extra_modules=None):
    cls = ggpm_b200.HierMPNEncoder if kind == "hier" else ggpm_b200.MotifEncoder
    model = cls(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), rnn, H, H, depthT, depthG, 0.0)
    model.load_state_dict({k: v for k, v in sd.items() if k in model.state_dict()}, strict=True)
    model = model.cuda()
    tree, graph = ingest.pack(tensorize.tensorize(mols)).to_device("cuda")
    masks, hooks = {}, []
    for site, mod in _relu_sites(model, kind).items():
        hooks.append(mod.register_forward_hook(
            lambda m, inp, out, site=site: masks.__setitem__(site, (inp[0][:, :H] > 0).cpu())))
    outs = model(tree, graph) if kind == "hier" else model(tree)
    for h in hooks:
        h.remove()
    loss(outs, model).backward()
    torch.cuda.synchronize()
    return [o.detach() for o in outs], {k: p.grad for k, p in model.named_parameters()}, masks


def _root_sum(outs, _):
    return outs[0].sum()


def _all_outputs(seed):
    """a fixed random cotangent on every output (so tree_encoder.W_o and every node table carry gradient)"""
    def loss(outs, _):
        gen = torch.Generator().manual_seed(seed)
        return sum((o * torch.randn(o.shape, generator=gen, dtype=torch.float64).to(o.dtype).to(o.device)).sum() for o in outs)
    return loss


def _compare(tag, kind, rnn, H, depth, B, seed, loss, min_atom_rows=0, oracle_device="cpu", grad_bar=None):
    mols = synth.random_batch(seed, B)
    sd = eo.random_state(kind, rnn, H, H, seed % 97)
    tree_np, graph_np = synth.tensorize(mols)
    if min_atom_rows:
        assert graph_np[1].shape[0] >= min_atom_rows, "batch too small to cross the plane-fed threshold: %d rows" % graph_np[1].shape[0]
    outs, grads, masks = _product(kind, rnn, H, depth, depth, sd, mols, loss)
    outs, grads = [o.cpu() for o in outs], {k: None if g is None else g.cpu() for k, g in grads.items()}
    torch.cuda.empty_cache()
    ref_outs, ref_grads, trace = _oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float64, loss,
                                         device=oracle_device)
    ref_outs32, ref_grads32, _ = _oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float32, loss,
                                         device=oracle_device)
    for name, a in zip(C.OUT_NAMES[kind][1:], outs[1:]):
        assert float(a[0].abs().max()) == 0.0, name                  # row 0 of every node table is exactly zero
    bad = C.compare_with_oracle(tag, C.OUT_NAMES[kind], outs, ref_outs, grads, ref_grads, ref_grads32, ref_outs32,
                                raise_on_bad=False, grad_bar=grad_bar)
    out_names = set(C.OUT_NAMES[kind])
    assert not [b for b in bad if b[0] in out_names], bad             # outputs are continuous across a kink: no excuse
    if bad or oracle_device != "cpu":          # sweep-size batches: always also record the kink-free comparison
        _kink_analysis(tag, kind, rnn, depth, sd, tree_np, graph_np, loss, masks, trace, outs, ref_outs, grads, bad,
                       oracle_device, grad_bar)


def _kink_analysis(tag, kind, rnn, depth, sd, tree_np, graph_np, loss, masks, trace, outs, ref_outs, grads, bad,
                   oracle_device="cpu", grad_bar=None):
    """gradient mismatches must be explained by ReLU units sitting on a kink: see the module docstring"""
    flips, worst_pre, n_units = 0, 0.0, 0
    for site, mask in masks.items():
        pre = trace[site]
        n_units += pre.numel()
        diff = mask != (pre > 0)
        diff[0] = False                                               # row 0 is masked to zero after the ReLU anyway
        n = int(diff.sum())
        if n:
            flips += n
            worst_pre = max(worst_pre, float(pre[diff].abs().max() / pre.abs().max()))
    max_flips = max(16, n_units // 250000)            # units within fp32 rounding of zero scale with the unit count
    C.record(tag, "kink:flipped_relu_units", flips, max_flips, relu_units=n_units, max_flipped_preactivation_rel=worst_pre,
             first_mismatch=str(bad[0]) if bad else None)
    assert (0 if not bad else 1) <= flips <= max_flips, ("gradient mismatch without a ReLU pattern difference" if flips == 0 else "too many flips", flips, bad)
    assert worst_pre <= 1e-5, ("a unit far from its kink took the other branch", worst_pre, bad)
    _, branch_grads, _ = _oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float64, loss,
                                 relu={k: v.double() for k, v in masks.items()}, device=oracle_device)
    # the reference arithmetic in fp32 on the SAME branch: its gap to fp64 is the measured fp32 noise of this tensor
    # (tests/_checks.py grad_tolerance), free of kink effects
    _, branch_grads32, _ = _oracle(kind, rnn, depth, depth, sd, tree_np, graph_np, torch.float32, loss,
                                   relu={k: v.float() for k, v in masks.items()}, device=oracle_device)
    C.compare_with_oracle(tag + ":on_product_relu_pattern", C.OUT_NAMES[kind], outs, ref_outs, grads, branch_grads,
                          branch_grads32, skip_outputs=True, grad_bar=grad_bar)


def test_hidden_500_depth_20_matches_fp64_oracle():
    _compare("full:hier_lstm_h500_d20_b32", "hier", "LSTM", 500, 20, 32, 50032, _root_sum)


def test_hidden_500_all_outputs_cotangent_matches_fp64_oracle():
    _compare("full:hier_lstm_h500_d20_b16_allout", "hier", "LSTM", 500, 20, 16, 50016, _all_outputs(3))


def test_plane_fed_forward_reached_naturally_matches_fp64_oracle():
    """480 molecules -> ~35.9k atom-layer message rows >= the library's 32 768-row threshold (api.cu sgather_min_rows):
    lstm_sgather + two TMA-fed forward GEMMs per round, exactly as at cfg3 / cfg4, with no GGPM_SGATHER_MIN_ROWS"""
    assert "GGPM_SGATHER_MIN_ROWS" not in os.environ
    _compare("full:hier_lstm_h250_d20_b480_planes", "hier", "LSTM", 250, 20, 480, 48000, _root_sum, min_atom_rows=32768)


def test_gru_hidden_250_depth_20_matches_fp64_oracle():
    _compare("full:hier_gru_h250_d20_b32", "hier", "GRU", 250, 20, 32, 25032, _all_outputs(5))


def test_motif_encoder_hidden_250_matches_fp64_oracle():
    _compare("full:motif_lstm_h250_d20_b256", "motif", "LSTM", 250, 20, 256, 25256, _all_outputs(7))


def test_sweep_batch_of_4096_molecules_matches_fp64_oracle_evaluated_on_the_gpu():
    """BASELINE config #4's batch (4096 molecules, hidden 250, depth 20/20: the default bench workload): ~306k atom-layer
    message rows, the K of the weight-gradient GEMMs.  This is the case that exposed the tensor-core accumulator's error
    growing with the length of the MMA chain (GGPM_SEG_CAP)."""
    _compare("full:hier_lstm_h250_d20_b4096_cfg4", "hier", "LSTM", 250, 20, 4096, 44096, _root_sum, min_atom_rows=250000,
             oracle_device="cuda")


def test_1024_molecules_hidden_500_matches_fp64_oracle_evaluated_on_the_gpu():
    """the width of BASELINE config #3 (hidden 500) at a batch the bench's cfg3s uses.  This case measured 1.09e-5 on four
    LSTM gradients until the glue GEMMs (3xTF32, post-ReLU inputs, K up to 2H) were cut into fp32-accumulated K-slices
    (profiles/r2_s14_glue_kslice_precision.txt): 8.3e-6 since."""
    _compare("full:hier_lstm_h500_d20_b1024_cfg3s", "hier", "LSTM", 500, 20, 1024, 31024, _root_sum, min_atom_rows=65536,
             oracle_device="cuda")


def test_cfg5_motif_encoder_plus_latent_head_matches_fp64_oracle():
    """BASELINE config #5, encoder side of PropertyVAE.forward (ggpm/property_vae.py:95-99): MotifEncoder ->
    rsample(R_mean, R_var) -> z, KL at hidden 250, latent 24, batch 256, same noise on both sides"""
    H, L, B, depth = 250, 24, 256, 20
    mols = synth.random_batch(55256, B)
    sd = eo.random_state("motif", "LSTM", H, H, 11)
    gen = torch.Generator().manual_seed(12)
    lat = {"R_mean.weight": torch.randn(L, H, generator=gen) / H ** 0.5, "R_mean.bias": torch.randn(L, generator=gen) * 0.1,
           "R_var.weight": torch.randn(L, H, generator=gen) / H ** 0.5, "R_var.bias": torch.randn(L, generator=gen) * 0.1}
    eps = torch.randn(B, L, generator=gen)
    cot = torch.randn(B, L, generator=gen)
    tree_np, graph_np = synth.tensorize(mols)

    model = ggpm_b200.MotifEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, depth, depth, 0.0)
    model.load_state_dict(sd, strict=True)
    model.R_mean, model.R_var = torch.nn.Linear(H, L), torch.nn.Linear(H, L)
    with torch.no_grad():
        for k, v in lat.items():
            dict(model.named_parameters())[k].copy_(v)
    model = model.cuda()
    tree, _ = ingest.pack((tensorize.tensorize(mols)[0], None)).to_device("cuda")
    root = model(tree)[0]
    z, kl = latent.rsample(root, model.R_mean, model.R_var, True, epsilon=eps.cuda())
    ((z * cot.cuda()).sum() + 0.7 * kl).backward()
    grads = {k: p.grad for k, p in model.named_parameters()}

    def oracle(dtype):
        torch.set_num_threads(os.cpu_count() or 1)
        leaves = {k: v.detach().clone().to(dtype).requires_grad_(True) for k, v in {**sd, **lat}.items()}
        r = eo.motif_encoder(leaves, "LSTM", depth, _as_long(tree_np))[0]
        zz, kk = lo.rsample(r, leaves["R_mean.weight"], leaves["R_mean.bias"], leaves["R_var.weight"],
                            leaves["R_var.bias"], eps.to(dtype), True)
        ((zz * cot.to(dtype)).sum() + 0.7 * kk).backward()
        return [r.detach(), zz.detach(), kk.detach()], {k: v.grad for k, v in leaves.items()}

    ref_outs, ref_grads = oracle(torch.float64)
    ref_outs32, ref_grads32 = oracle(torch.float32)
    C.compare_with_oracle("full:cfg5_motif_latent_h250_l24_b256", ("hroot", "z", "kl"), [root.detach(), z.detach(), kl.detach()],
                          ref_outs, grads, ref_grads, ref_grads32, ref_outs32)
