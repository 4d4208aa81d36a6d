"""Drop-in boundary through the reference's OWN model wrappers: the unmodified
`HierPropertyVAE` / `HierVAE` / `PropertyVAE` (ggpm/property_vae.py:11-20,76-92; ggpm/hpgnn.py:50-58) are built with the
three rebinding lines of INTEGRATION.md section 2, and then used exactly as the reference's scripts use them:

  * `tie_embedding` hands our encoder the decoder's `nn.Sequential(Embedding, Dropout)` objects (ggpm/encoder.py:92-94);
  * a `state_dict` produced by the UNPATCHED reference model loads strictly (vae_train.py:55-56, reconstruct.py:53);
  * `copy_encoder` (ggpm/nnutils.py:40-62) filters by key and shape and loads strictly;
  * the wrapper's own encoder line — `make_cuda(tensors)` then `self.encoder(tree_tensors, graph_tensors)`
    (ggpm/property_vae.py:48-50) — gives `root_vecs` and encoder gradients equal to the fp64 oracle's (pinned to the
    reference at 1e-12) and to the unpatched reference model's own fp32 result; the tied E_c / E_i, which live in the
    decoder's module tree, receive the encoder's gradient.

The reference modules come from /root/reference or the staged byte-identical copy oracle/_ref (oracle/make_ref.py), so the
test runs on the CPU emulator here and on the B200 (-m gpu) there.
"""
import os
import sys
import types

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _checks as C                                                                     # noqa: E402
import _emu_lib                                                                        # noqa: E402
import ggpm_b200                                                                       # noqa: E402
from ggpm_b200 import _lib, nnutils, synth                                              # noqa: E402
from oracle import encoder_oracle as eo                                                # noqa: E402
from oracle import ref_loader as rl                                                    # noqa: E402

pytestmark = pytest.mark.skipif(not rl.reference_available(), reason="no copy of the reference reachable")
NM, NA, H, L = 50, 60, 32, 8


def _args(dropout=0.0):
    ref = rl.load_reference()
    return types.SimpleNamespace(vocab=rl.FakePairVocab(NM, NA), atom_vocab=ref.vocab.common_atom_vocab, rnn_type="LSTM",
                                 embed_size=H, hidden_size=H, depthT=3, depthG=4, dropout=dropout, latent_size=L,
                                 diterT=1, diterG=3, tie_embedding=True)


class _Rebound:
    """the three lines of INTEGRATION.md section 2, undone on exit"""

    def __enter__(self):
        self.pv, self.hp = rl.load_reference_module("property_vae"), rl.load_reference_module("hpgnn")
        self.saved = (self.pv.HierMPNEncoder, self.pv.MotifEncoder, self.hp.HierMPNEncoder)
        self.hp.HierMPNEncoder = ggpm_b200.HierMPNEncoder                 # HierVAE            (ggpm/hpgnn.py:50)
        self.pv.HierMPNEncoder = ggpm_b200.HierMPNEncoder                 # HierPropertyVAE …  (ggpm/property_vae.py:14)
        self.pv.MotifEncoder = ggpm_b200.MotifEncoder                     # PropertyVAE …      (ggpm/property_vae.py:79)
        return self

    def __exit__(self, *exc):
        self.pv.HierMPNEncoder, self.pv.MotifEncoder, self.hp.HierMPNEncoder = self.saved


def _init(model, seed):
    torch.manual_seed(seed)
    for p in model.parameters():                                           # vae_train.py:49-53, but non-zero biases
        if p.dim() == 1:
            torch.nn.init.normal_(p, std=0.1)
        else:
            torch.nn.init.xavier_normal_(p)


def _batch(seed=31, n=6):
    mols = synth.random_batch(seed, n, n_motif_vocab=NM, n_attach_vocab=NA, mean_motifs=6.0, sd_motifs=1.5)
    return synth.tensorize(mols)


def _wrapper_case(kind, device, tmp_path):
    ref = rl.load_reference()
    pv, hp = rl.load_reference_module("property_vae"), rl.load_reference_module("hpgnn")
    cls_of = lambda: {"hier-prop": pv.HierPropertyVAE, "hier": hp.HierVAE, "prop": pv.PropertyVAE}[kind]
    args = _args()

    # 1. the unpatched reference model (kept on the host: its one-hot tables are plain tensors placed at construction,
    #    ggpm/encoder.py:74-77): its state_dict is what checkpoints hold
    ref.nnutils.device = torch.device("cpu")
    ref_model = cls_of()(args)
    ref.nnutils.device = torch.device(device)          # the reference's global device (ggpm/nnutils.py:9-10); one process per GPU
    assert type(ref_model.encoder).__module__ == "ggpm.encoder"
    _init(ref_model, 5)
    ref_sd = {k: v.detach().clone() for k, v in ref_model.state_dict().items()}
    path = os.path.join(str(tmp_path), "model.ckpt")
    torch.save(ref_sd, path)

    with _Rebound():
        # 2. same constructor call, our encoder inside, embeddings tied to the decoder's modules
        model = cls_of()(args)
        assert isinstance(model.encoder, ggpm_b200.MotifEncoder if kind == "prop" else ggpm_b200.HierMPNEncoder)
        assert model.encoder.E_c is model.decoder.hmpn.E_c and model.encoder.E_i is model.decoder.hmpn.E_i
        assert type(model.decoder.hmpn).__module__ == "ggpm.encoder"            # the decoder keeps the reference's own Inc* classes
        assert list(model.state_dict().keys()) == list(ref_sd.keys())
        assert [tuple(v.shape) for v in model.state_dict().values()] == [tuple(v.shape) for v in ref_sd.values()]
        assert [n for n, _ in model.named_parameters()] == [n for n, _ in ref_model.named_parameters()]   # Adam / init order
        # 3. strict load of the reference-made checkpoint
        model.load_state_dict(torch.load(path, map_location="cpu"), strict=True)
        # 4. copy_encoder: (to-be-copied model, scratch model to load the file into, path)
        fresh = cls_of()(args)
        _init(fresh, 99)
        ref.nnutils.copy_encoder(fresh, cls_of()(args), path)
        for (k, a), (_, b) in zip(fresh.encoder.state_dict().items(), model.encoder.state_dict().items()):
            assert torch.equal(a, b), k
        model = ref.nnutils.to_cuda(model)                                      # vae_train.py:36-37
        model.train()

        # 5. the wrapper's encoder line (ggpm/property_vae.py:48-50 / 101-104, ggpm/hpgnn.py:58-60)
        tree_np, graph_np = _batch()
        tree_tensors, graph_tensors = pv.make_cuda((tree_np, graph_np))
        assert all(t.dtype == torch.long for t in tree_tensors[:-1])
        if kind == "prop":
            root_vecs, _ = model.encoder(tree_tensors)
        else:
            root_vecs, _, _, _ = model.encoder(tree_tensors, graph_tensors)
        z, kl = (model.rsample(root_vecs, False) if kind == "prop" else model.rsample(root_vecs, model.R_mean, model.R_var, False))
        (root_vecs.sum() + z.sum() + kl).backward()
        got = {k: p.grad for k, p in model.named_parameters()}

    # ---- expected: the fp64 oracle on the encoder's part of the checkpoint, and the unpatched reference model itself
    enc_sd = {k[len("encoder."):]: v for k, v in ref_sd.items() if k.startswith("encoder.")}
    as_long = lambda ts: [torch.from_numpy(np.asarray(t)).long() if not isinstance(t, list) else t for t in ts]
    leaves = {k: v.double().requires_grad_(True) for k, v in enc_sd.items()}
    if kind == "prop":
        o_root = eo.motif_encoder(leaves, "LSTM", args.depthT, as_long(tree_np))[0]
    else:
        o_root = eo.hier_encoder(leaves, "LSTM", args.depthT, args.depthG, as_long(tree_np), as_long(graph_np))[0]
    Wm, bm, Wv, bv = (ref_sd[k].double() for k in ("R_mean.weight", "R_mean.bias", "R_var.weight", "R_var.bias"))
    zm = o_root @ Wm.t() + bm
    zlv = -torch.abs(o_root @ Wv.t() + bv)
    o_kl = -0.5 * torch.sum(1.0 + zlv - zm * zm - torch.exp(zlv)) / o_root.shape[0]
    (o_root.sum() + zm.sum() + o_kl).backward()
    assert C.rel(root_vecs, o_root) <= C.TOL_OUT, C.rel(root_vecs, o_root)
    for k, v in leaves.items():
        g = got["encoder." + k] if "encoder." + k in got else got.get("decoder.hmpn." + k)
        if v.grad is None or float(v.grad.abs().max()) == 0.0:
            assert g is None or float(g.abs().max()) == 0.0, k
        else:
            assert g is not None and C.rel(g, v.grad) <= C.TOL_GRAD, (k, None if g is None else C.rel(g, v.grad))
    # the tied tables are registered under the decoder (the encoder's own copies were replaced): they carry the gradient
    assert model.decoder.hmpn.E_c[0].weight.grad is not None and model.decoder.hmpn.E_i[0].weight.grad is not None
    assert C.rel(model.decoder.hmpn.E_c[0].weight.grad, leaves["E_c.0.weight"].grad) <= C.TOL_GRAD

    # the unpatched reference model, fp32, on the host: same call, same numbers up to fp32 rounding
    ref.nnutils.device = torch.device("cpu")
    ref_model = ref_model.cpu().train()
    t_ref, g_ref = pv.make_cuda((tree_np, graph_np))
    r_root = ref_model.encoder(t_ref)[0] if kind == "prop" else ref_model.encoder(t_ref, g_ref)[0]
    assert C.rel(root_vecs, r_root) <= 2e-5, C.rel(root_vecs, r_root)


@pytest.fixture()
def emu():
    lib = _emu_lib.EmuLib()
    _lib._install_test_backend(lib)
    nnutils.set_default_device("cpu")
    yield lib
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)
    rl.load_reference().nnutils.device = torch.device("cpu")


@pytest.mark.parametrize("kind", ["hier-prop", "hier", "prop"])
def test_reference_wrappers_around_the_drop_in_encoder_on_the_emulator(emu, kind, tmp_path):
    _wrapper_case(kind, "cpu", tmp_path)
    assert "lstm_fwd" in emu.calls and "lstm_bwd" in emu.calls


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["hier-prop", "hier", "prop"])
def test_reference_wrappers_around_the_drop_in_encoder_on_b200(kind, tmp_path):
    assert _lib._TEST_BACKEND is None
    before = _lib.launch_count()
    try:
        _wrapper_case(kind, "cuda:0", tmp_path)
    finally:
        rl.load_reference().nnutils.device = torch.device("cpu")
    assert _lib.launch_count() > before
