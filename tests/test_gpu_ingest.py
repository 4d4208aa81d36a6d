"""GPU: batch ingest — the host-built CSR / reverse CSR of ingest.PackedBatch against the device builders of
csrc/csr.cu (bit-exact), and the encoder driven through a packed batch against the reference golden vectors."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _checks as C                                                                     # noqa: E402
from ggpm_b200 import _lib, functional as Fn, ingest, nnutils, synth                   # noqa: E402

pytestmark = pytest.mark.gpu


def _same_csr(host: Fn.CSR, dev: Fn.CSR):
    nnz = int(dev.row_ptr[-1].item())
    assert host.rows == dev.rows and host.n_src == dev.n_src
    assert torch.equal(host.row_ptr, dev.row_ptr)
    assert torch.equal(host.col[:nnz], dev.col[:nnz])
    assert torch.equal(host.T.row_ptr, dev.T.row_ptr)
    assert torch.equal(host.T.col[:nnz], dev.T.col[:nnz])
    assert torch.equal(host.T.pos[:nnz], dev.T.pos[:nnz])


@pytest.mark.parametrize("seed,n", [(11, 1), (12, 20), (13, 256)])
def test_host_built_index_structures_equal_device_built(seed, n):
    tree, graph = synth.tensorize(synth.random_batch(seed, n))
    pt, pg = ingest.pack((tree, graph)).to_device("cuda")
    dt, dg = nnutils.make_cuda((tree, graph))
    n_tm, n_gm, n_gn, n_tn = tree[1].shape[0], graph[1].shape[0], graph[0].shape[0], tree[0].shape[0]
    for packed, plain, n_src in ((pt[2], dt[2], n_tm), (pt[3], dt[3], n_tm), (pt[4], dt[4], n_gn),
                                 (pg[2], dg[2], n_gm), (pg[3], dg[3], n_gm)):
        assert packed.dtype == torch.int32 and torch.equal(packed.long(), plain)
        _same_csr(packed._ggpm_csr, Fn.csr_from_padded(plain, n_src))
    for packed, plain, n_nodes in ((pt[1], dt[1], n_tn), (pg[1], dg[1], n_gn)):
        dev_by_src = Fn.csr_from_padded(plain[:, 0].to(torch.int32).contiguous().view(-1, 1), n_nodes).T
        host_by_src = packed._ggpm_by_src
        nnz = int(dev_by_src.row_ptr[-1].item())
        assert torch.equal(host_by_src.row_ptr, dev_by_src.row_ptr)
        assert torch.equal(host_by_src.col[:nnz], dev_by_src.col[:nnz])


@pytest.mark.parametrize("name", ["hier_lstm_small", "hier_gru_small", "motif_lstm_small", "hier_lstm_deep"])
def test_packed_batch_path_matches_reference_golden(name):
    C.check_forward_backward_golden(name, "cuda", packed=True)


def test_packed_path_launches_no_index_kernels_and_matches_plain_path():
    tree, graph = synth.tensorize(synth.random_batch(21, 24))
    import ggpm_b200
    torch.manual_seed(0)
    model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", 128, 128, 4, 4, 0.0).cuda()
    pb = ingest.pack((tree, graph))
    outs = {}
    for kind in ("plain", "packed"):
        t, g = nnutils.make_cuda((tree, graph)) if kind == "plain" else pb.to_device("cuda")
        model.zero_grad(set_to_none=True)
        n0 = _lib.launch_count()
        o = model(t, g)
        o[0].sum().backward()
        torch.cuda.synchronize()
        outs[kind] = (_lib.launch_count() - n0, [x.detach().clone() for x in o],
                      {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None})
    assert outs["packed"][0] < outs["plain"][0]                  # the CSR / transpose kernels are gone
    for a, b in zip(outs["plain"][1], outs["packed"][1]):
        assert C.rel(a, b) <= 1e-6                               # same kernels on identical index arrays (fp32 atomics aside)
    for k, g in outs["plain"][2].items():
        assert C.rel(outs["packed"][2][k], g) <= 1e-5, k


def test_device_loader_overlapped_copies_deliver_the_same_tensors():
    pbs = [ingest.pack(synth.tensorize(synth.random_batch(60 + i, 8 + 5 * i))) for i in range(5)]
    x = torch.randn(4096, 4096, device="cuda")
    for depth in (1, 2, 3):
        for pb, (tree, graph) in zip(pbs, ingest.DeviceLoader(iter(pbs), "cuda", depth=depth)):
            y = x @ x                                   # keep the compute stream busy while the next copy runs
            ref_tree, ref_graph = pb.to_device("cuda")
            for a, b in zip(tree[:5] + graph[:4], ref_tree[:5] + ref_graph[:4]):
                assert torch.equal(a, b)
            assert torch.equal(tree[3]._ggpm_csr.T.col, ref_tree[3]._ggpm_csr.T.col)
            assert torch.equal(tree[5].roots_dev, ref_tree[5].roots_dev)
    torch.cuda.synchronize()
