"""CPU, world_size 2 over gloo: the data-parallel path of SURVEY 8(e) — batches sharded by molecule with no
collective on the encoder path, one flat-bucket all-reduce of the encoder-weight gradients.  The encoder runs on
the C-ABI emulator (tests/_emu_lib.py); what is tested is the sharding + bucket logic."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _model(H=32, depth=3):
    import ggpm_b200
    from ggpm_b200 import synth
    torch.manual_seed(3)
    m = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(50, 60), synth.SizeOnlyAtomVocab(), "LSTM", H, H, depth, depth, 0.0)
    for p in m.parameters():
        torch.nn.init.normal_(p, std=0.1)
    return m


def _grads(model, mols):
    from ggpm_b200 import synth
    tree, graph = synth.tensorize(mols)
    as_long = lambda ts: [torch.from_numpy(np.asarray(t)).long() if not isinstance(t, list) else t for t in ts]
    model.zero_grad(set_to_none=True)
    model(as_long(tree), as_long(graph))[0].sum().backward()


def _worker(rank, world, port, weighted, out_dir):
    import _emu_lib
    from ggpm_b200 import _lib, nnutils, synth
    from ggpm_b200.dist import GradBucket, shard_batches
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    _lib._install_test_backend(_emu_lib.EmuLib())
    nnutils.set_default_device("cpu")
    mols = synth.random_batch(77, 6, n_motif_vocab=50, n_attach_vocab=60)
    mine = mols[:2] if (weighted and rank == 0) else mols[2:] if weighted else [mols[i] for i in shard_batches(6, rank, world)]
    model = _model()
    if weighted:
        bucket = GradBucket(model.parameters())
        _grads(model, mine)
        assert model.tree_encoder.W_o[0].weight.grad is None        # root-only loss: contributes zeros to the bucket
        bucket.all_reduce_mean(weight=float(len(mine)))
    else:
        # two groups (the atom layer last), early group launched from autograd hooks: the first step learns which
        # parameters receive gradients and reduces everything at the end, the second launches the early group while
        # the backward pass is still running
        bucket = GradBucket.for_encoder(model)
        assert len(bucket.bounds) == 2
        for step in range(2):
            bucket.arm()
            _grads(model, mine)
            early_launched = bucket._launched[0]
            bucket.all_reduce_mean()
            assert early_launched == (step == 1), (step, early_launched)
        _grads(model, mine)                                          # an un-armed backward pass launches nothing
        assert not any(bucket._launched) and not bucket._works
        bucket.arm()
        _grads(model, mine)
        bucket.all_reduce_mean()
    torch.save({k: p.grad.clone() for k, p in model.named_parameters()}, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("weighted", [False, True])
def test_two_rank_gradient_allreduce_equals_full_batch(tmp_path, weighted):
    import _emu_lib
    from ggpm_b200 import _lib, nnutils, synth
    mp.spawn(_worker, args=(2, _free_port(), weighted, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r)) for r in (0, 1))
    for k in r0:
        assert torch.equal(r0[k], r1[k]), k                           # every rank holds the same reduced gradient
    # single process, whole batch: loss = sum over molecules, so the reduced gradient is full / divisor
    _lib._install_test_backend(_emu_lib.EmuLib())
    nnutils.set_default_device("cpu")
    try:
        model = _model()
        mols = synth.random_batch(77, 6, n_motif_vocab=50, n_attach_vocab=60)
        if weighted:      # ranks held molecules [0:2] and [2:6]: sum_r n_r g_r / sum_r n_r
            parts = []
            for part in (mols[:2], mols[2:]):
                _grads(model, part)
                parts.append({k: (None if p.grad is None else p.grad.clone()) for k, p in model.named_parameters()})
            expect = {k: (None if parts[0][k] is None else (2.0 * parts[0][k] + 4.0 * parts[1][k]) / 6.0) for k in parts[0]}
        else:             # equal shards: loss = sum over molecules, so the mean over 2 ranks is full / 2
            _grads(model, mols)
            expect = {k: (None if p.grad is None else p.grad / 2.0) for k, p in model.named_parameters()}
        for k, e in expect.items():
            if e is None:
                assert float(r0[k].abs().max()) == 0.0, k
            else:
                assert torch.allclose(r0[k], e, rtol=2e-4, atol=1e-6), k
    finally:
        _lib._install_test_backend(None)
        nnutils.set_default_device(None)


def test_bucket_pack_unpack_single_process():
    from ggpm_b200.dist import GradBucket, shard_batches
    assert shard_batches(7, 1, 3) == [1, 4] and shard_batches(2, 3, 4) == []
    a, b, c = (torch.nn.Parameter(torch.randn(*s)) for s in ((3, 4), (5,), (2, 2)))
    a.grad, c.grad = torch.randn(3, 4), torch.randn(2, 2)
    ga, gc = a.grad.clone(), c.grad.clone()
    bk = GradBucket([a, b, c, a])                                     # tied parameter listed twice: counted once
    assert bk.numel == 12 + 5 + 4
    flat = bk.pack()
    assert torch.equal(flat[:12].view(3, 4), ga) and float(flat[12:17].abs().max()) == 0 and torch.equal(flat[17:].view(2, 2), gc)
    bk.all_reduce_mean()                                              # no process group: identity
    assert torch.equal(a.grad, ga) and torch.equal(c.grad, gc) and float(b.grad.abs().max()) == 0
    bk.all_reduce_mean(weight=3.0)                                    # weight cancels on one rank
    assert torch.allclose(a.grad, ga) and torch.allclose(c.grad, gc)
