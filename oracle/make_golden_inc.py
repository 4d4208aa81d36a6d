"""TEST INFRASTRUCTURE ONLY — generate tests/golden_inc/*.npz by running the UNMODIFIED reference's incremental
encoders (ggpm/encoder.py:160-249, 344-394; ggpm/rnn.py:52-59, 110-121; ggpm/nnutils.py:124-128).

Run in the build container (where /root/reference exists):

    python -m oracle.make_golden_inc

For every case: a seeded synthetic batch and its decode schedule (oracle/inc_oracle.py: decode_schedule), the
reference ``IncHierMPNEncoder`` / ``IncEncoder`` in fp64 with a seeded state_dict driven by ``run_decode`` (the
decoder's teacher-forced loop restated around the encoder), seeded cotangents on every collected output, backward.
Asserts oracle/inc_oracle.py agrees with the reference to <= 1e-12 on every output, every parameter gradient and
the gradient of the initial root vectors — the oracle's parity pin — and writes outputs and gradients (fp64) to
tests/golden_inc/<case>.npz.  Inputs are not stored: batch, schedule, weights and cotangents are regenerated from
the seeds in the case record (a checksum of the index tensors and of the schedule is stored and re-checked).
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ggpm_b200 import synth                                    # noqa: E402
from oracle import encoder_oracle as eo                        # noqa: E402
from oracle import inc_oracle as io                            # noqa: E402
from oracle.ref_loader import FakePairVocab, IdAtomVocab, load_reference  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden_inc")

CASES = [
    dict(name="inc_hier_lstm", kind="hier", rnn="LSTM", hidden=24, depthT=2, depthG=3, batch=4, data_seed=21,
         weight_seed=201, vocab=(40, 60), mean_motifs=4.0),
    dict(name="inc_hier_gru", kind="hier", rnn="GRU", hidden=24, depthT=2, depthG=3, batch=4, data_seed=22,
         weight_seed=202, vocab=(40, 60), mean_motifs=4.0),
    dict(name="inc_motif_lstm", kind="motif", rnn="LSTM", hidden=20, depthT=3, depthG=0, batch=5, data_seed=23,
         weight_seed=203, vocab=(40, 60), mean_motifs=5.0),
    dict(name="inc_motif_gru", kind="motif", rnn="GRU", hidden=20, depthT=3, depthG=0, batch=5, data_seed=24,
         weight_seed=204, vocab=(40, 60), mean_motifs=5.0),
    # the decoder's shipped iteration counts (configs/*.json: diterT = 1, diterG = 5) at an unaligned width
    dict(name="inc_hier_lstm_diter", kind="hier", rnn="LSTM", hidden=34, depthT=1, depthG=5, batch=5, data_seed=25,
         weight_seed=205, vocab=(40, 60), mean_motifs=5.0),
]


def build_inputs(case, device="cpu"):
    """(mols, tree_tensors, graph_tensors, schedule, init_vecs fp64 [B, H]) regenerated from the case's seeds"""
    nm, na = case["vocab"]
    mols = synth.random_batch(case["data_seed"], case["batch"], n_motif_vocab=nm, n_attach_vocab=na,
                              mean_motifs=case["mean_motifs"], sd_motifs=1.0)
    tree_np, graph_np = synth.tensorize(mols)
    sched = io.decode_schedule(mols, tree_np, graph_np)
    as_long = lambda ts: [torch.from_numpy(np.ascontiguousarray(t)).long().to(device) for t in ts[:-1]] + [ts[-1]]
    gen = torch.Generator().manual_seed(case["weight_seed"] + 3)
    init = torch.randn(case["batch"], case["hidden"], generator=gen, dtype=torch.float64) * 0.5
    return mols, as_long(tree_np), as_long(graph_np), sched, init


def inputs_digest(tree_t, graph_t, sched) -> str:
    h = hashlib.sha256()
    for t in list(tree_t[:-1]) + list(graph_t[:-1]):
        h.update(np.ascontiguousarray(t.cpu().numpy()).tobytes())
    h.update(json.dumps([tree_t[-1], graph_t[-1]]).encode())
    h.update(json.dumps(sched, sort_keys=True).encode())
    return h.hexdigest()


def inc_state(case, dtype=torch.float32):
    """reference-keyed state_dict of the incremental encoder: the encoder's keys without W_root (deleted by the
    Inc constructors, ggpm/encoder.py:195, 350)"""
    nm, na = case["vocab"]
    sd = eo.random_state(case["kind"], case["rnn"], case["hidden"], case["hidden"], case["weight_seed"],
                         n_motif=nm, n_attach=na)
    return {k: v.to(dtype) for k, v in sd.items() if not k.startswith("W_root.")}


def rel_err(a, b):
    a, b = a.detach(), b.detach()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-300)


def run_case(ref, case):
    nm, na = case["vocab"]
    hier = case["kind"] == "hier"
    mols, tree_t, graph_t, sched, init = build_inputs(case)
    sd64 = inc_state(case, torch.float64)

    torch.set_default_dtype(torch.float64)
    try:
        cls = ref.encoder.IncHierMPNEncoder if hier else ref.encoder.IncEncoder
        model = cls(FakePairVocab(nm, na), IdAtomVocab(), case["rnn"], case["hidden"], case["hidden"], case["depthT"],
                    case["depthG"], 0.0)
        assert list(model.state_dict().keys()) == list(sd64.keys()), "state_dict key order differs"
        model.load_state_dict(sd64, strict=True)
        model.double().eval()
        init_r = init.clone().requires_grad_(True)
        outs = io.run_decode(model, hier, tree_t, graph_t, sched, init_r)
        cots = io.cotangents(outs, case["weight_seed"] + 7)
        io.scalar_loss(outs, cots).backward()
        grads = {k: (p.grad.clone() if p.grad is not None else torch.zeros_like(p)) for k, p in model.named_parameters()}
        grads["__init_vecs"] = init_r.grad.clone()
    finally:
        torch.set_default_dtype(torch.float32)

    # --- the oracle must agree with the reference (parity pin) ------------------------------------
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd64.items()}
    orc = io.IncHierOracle(sdo, case["rnn"], case["depthT"], case["depthG"]) if hier else \
        io.IncMotifOracle(sdo, case["rnn"], case["depthT"])
    init_o = init.clone().requires_grad_(True)
    o_outs = io.run_decode(orc, hier, tree_t, graph_t, sched, init_o)
    assert len(o_outs) == len(outs)
    worst = max(rel_err(a, b) for a, b in zip(o_outs, outs) if b.numel())
    io.scalar_loss(o_outs, cots).backward()
    o_grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in sdo.items()}
    o_grads["__init_vecs"] = init_o.grad
    for k, g in grads.items():
        if float(g.abs().max()) > 0:
            worst = max(worst, rel_err(o_grads[k], g))
        else:
            assert float(o_grads[k].abs().max()) == 0.0, k
    assert worst <= 1e-12, (case["name"], worst)

    # --- the reference in fp32: its own fp32-vs-fp64 headroom on this path --------------------------
    model32 = cls(FakePairVocab(nm, na), IdAtomVocab(), case["rnn"], case["hidden"], case["hidden"], case["depthT"],
                  case["depthG"], 0.0)
    model32.load_state_dict(inc_state(case), strict=True)
    model32.eval()
    with torch.no_grad():
        outs32 = io.run_decode(model32, hier, tree_t, graph_t, sched, init.float())
    fp32_gap = max(rel_err(a.double(), b) for a, b in zip(outs32, outs) if b.numel())

    blob = {"meta": np.frombuffer(json.dumps(dict(case, n_outputs=len(outs), n_steps=len(sched["steps"]),
                                                  digest=inputs_digest(tree_t, graph_t, sched), fp32_gap=fp32_gap,
                                                  oracle_vs_ref=worst)).encode(), dtype=np.uint8)}
    for i, o in enumerate(outs):
        blob["out_%03d" % i] = o.detach().numpy()
    for k, g in grads.items():
        blob["grad_" + k] = g.numpy()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, case["name"] + ".npz")
    np.savez_compressed(path, **blob)
    print("%-22s steps %3d outputs %3d  oracle-vs-ref %.2e  ref fp32 gap %.2e  %6.1f KB"
          % (case["name"], len(sched["steps"]), len(outs), worst, fp32_gap, os.path.getsize(path) / 1024))


def main():
    ref = load_reference()
    for case in CASES:
        run_case(ref, case)


if __name__ == "__main__":
    main()
