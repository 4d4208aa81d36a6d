"""CPU: batch ingest (ggpm_b200/ingest.py) — the host CSR builders against the index oracle, bit for bit, on seeded
synthetic batches and hand-made edge cases; PackedBatch layout, wire format round trip, reference-layout views."""
import io
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import load_golden                                                       # noqa: E402
from ggpm_b200 import ingest, synth                                                    # noqa: E402
from oracle import tensorize_oracle as to                                              # noqa: E402


def _edge_cases():
    rng = np.random.default_rng(5)
    yield np.zeros((1, 1), dtype=np.int32)                                   # only the padding row
    yield np.zeros((7, 3), dtype=np.int32)                                   # no neighbours at all
    yield np.array([[0, 0, 0], [2, 0, 3], [0, 1, 0], [3, 3, 0]], dtype=np.int32)   # mid-row zeros, duplicates
    yield np.arange(12, dtype=np.int32).reshape(4, 3) % 4                    # dense, repeated ids
    big = rng.integers(0, 500, size=(500, 6)).astype(np.int32)
    big[rng.random(big.shape) < 0.6] = 0
    yield big
    yield rng.integers(1, 40, size=(40,)).astype(np.int32)                   # 1-D (the per-message source column)


@pytest.mark.parametrize("case", range(6))
def test_host_csr_matches_oracle(case):
    idx = list(_edge_cases())[case]
    rows = idx.shape[0]
    rp, col = ingest.csr_from_padded_host(idx)
    orp, ocol = to.csr_from_padded_oracle(idx.reshape(rows, -1))
    assert rp.dtype == np.int32 and col.dtype == np.int32
    assert np.array_equal(rp, orp) and np.array_equal(col, ocol)
    n_src = max(rows, int(idx.max()) + 1)
    tp, trow, tpos = ingest.csr_transpose_host(rp, col, n_src)
    otp, otrow, otpos = to.csr_transpose_oracle(orp, ocol, n_src)
    assert np.array_equal(tp, otp) and np.array_equal(trow, otrow) and np.array_equal(tpos, otpos)
    # every (e, n) pair is listed once on each side
    for n in range(n_src):
        for j in range(tp[n], tp[n + 1]):
            assert col[tpos[j]] == n and rp[trow[j]] <= tpos[j] < rp[trow[j] + 1]


def test_transpose_rejects_out_of_range():
    rp, col = ingest.csr_from_padded_host(np.array([[0, 0], [5, 0]], dtype=np.int32))
    with pytest.raises(IndexError):
        ingest.csr_transpose_host(rp, col, 3)
    with pytest.raises(IndexError):
        ingest.PackedBatch.from_tensors(*_bad_batch())


def _bad_batch():
    tree, graph = synth.tensorize(synth.random_batch(3, 4))
    tree = list(tree)
    bg = tree[3].copy()
    bg[1, 0] = bg.shape[0] + 7
    tree[3] = bg
    return tree, graph


@pytest.mark.parametrize("seed,n", [(1, 1), (2, 5), (3, 32)])
def test_packed_batch_views_and_structures(seed, n):
    tree, graph = synth.tensorize(synth.random_batch(seed, n))
    pb = ingest.PackedBatch.from_tensors(tree, graph, pin=False)
    assert pb.host.dtype == torch.int32 and pb.nbytes == pb.host.numel() * 4
    for off, _ in pb.entries.values():
        assert off % 4 == 0                                                   # 16-byte aligned segments
    dt, dg = pb.to_device("cpu")
    for a, b in zip(tree[:5], dt[:5]):
        assert b.dtype == torch.int32 and np.array_equal(np.asarray(a), b.numpy())
    for a, b in zip(graph[:4], dg[:4]):
        assert np.array_equal(np.asarray(a), b.numpy())
    assert list(dt[5]) == [tuple(s) for s in tree[5]] and list(dg[4]) == [tuple(s) for s in graph[4]]
    assert dt[5].roots_dev.tolist() == [st for st, _ in tree[5]]
    n_tm, n_gm, n_gn = tree[1].shape[0], graph[1].shape[0], graph[0].shape[0]
    for t, idx, n_src in ((dt[2], tree[2], n_tm), (dt[3], tree[3], n_tm), (dt[4], tree[4], n_gn),
                          (dg[2], graph[2], n_gm), (dg[3], graph[3], n_gm)):
        csr = t._ggpm_csr
        orp, ocol = to.csr_from_padded_oracle(np.asarray(idx))
        nnz = int(orp[-1])
        assert csr.rows == idx.shape[0] and csr.n_src == n_src
        assert np.array_equal(csr.row_ptr.numpy(), orp) and np.array_equal(csr.col.numpy()[:nnz], ocol)
        otp, otrow, otpos = to.csr_transpose_oracle(orp, ocol, n_src)
        assert np.array_equal(csr.T.row_ptr.numpy(), otp)
        assert np.array_equal(csr.T.col.numpy()[:nnz], otrow) and np.array_equal(csr.T.pos.numpy()[:nnz], otpos)
        assert csr.T.T is csr
    src, k1, k2 = dg[1]._ggpm_cols
    assert np.array_equal(src.numpy(), graph[1][:, 0]) and np.array_equal(k1.numpy(), graph[1][:, 2])
    assert np.array_equal(k2.numpy(), graph[1][:, 3] + 4)
    by = dg[1]._ggpm_by_src                                                  # node -> messages it emits, ascending
    for v in range(n_gn):
        mine = by.col.numpy()[by.row_ptr[v]:by.row_ptr[v + 1]]
        assert mine.tolist() == [e for e in range(n_gm) if graph[1][e, 0] == v and v != 0]


def test_tree_only_batch_and_wire_format_roundtrip():
    meta, z, tree, graph = load_golden("hier_lstm_small")
    pb = ingest.pack((tree, graph), pin=False)
    f = io.BytesIO()
    pb.save(f)
    f.seek(0)
    pb2 = ingest.PackedBatch.load(f, pin=False)
    assert torch.equal(pb.host, pb2.host) and pb.entries == pb2.entries
    assert pb.tree_scope == pb2.tree_scope and pb.graph_scope == pb2.graph_scope
    f.seek(0)
    with pytest.raises(ValueError):
        ingest.PackedBatch.load(io.BytesIO(f.read()[:-8]), pin=False)
    with pytest.raises(ValueError):
        ingest.PackedBatch.load(io.BytesIO(b"not a batch at all"), pin=False)
    t_only, none = ingest.PackedBatch.from_tensors(tree, None, pin=False).to_device("cpu")
    assert none is None and not hasattr(t_only[4], "_ggpm_csr") and hasattr(t_only[3], "_ggpm_csr")


def test_device_loader_yields_every_batch_in_order():
    pbs = [ingest.pack(synth.tensorize(synth.random_batch(40 + i, 3 + i)), pin=False) for i in range(4)]
    got = list(ingest.DeviceLoader(iter(pbs), "cpu", depth=2))
    assert len(got) == 4
    for pb, (tree, graph) in zip(pbs, got):
        ref_tree, ref_graph = pb.to_device("cpu")
        for a, b in zip(tree[:5] + graph[:4], ref_tree[:5] + ref_graph[:4]):
            assert torch.equal(a, b)
        assert hasattr(tree[3], "_ggpm_csr") and list(tree[5]) == list(ref_tree[5])
    assert list(ingest.DeviceLoader(iter([]), "cpu")) == []
