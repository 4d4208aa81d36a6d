"""B200: the incremental encoders through the CUDA library — golden vectors of the unmodified
reference (tests/golden_inc), the fp64 oracle at the shipped width (H = 250, diterT = 1, diterG = 5), index_scatter and
the sparse cells with mid-row zero slots."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _inc_checks as IC                                               # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _device():
    from ggpm_b200 import _lib, nnutils
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)
    yield


@pytest.mark.parametrize("name", IC.inc_golden_names())
def test_incremental_encoders_match_reference_golden(name):
    from ggpm_b200 import _lib
    n0 = _lib.launch_count()
    IC.check_golden(name, DEV)
    assert _lib.launch_count() > n0, "the CUDA library did not run"


def test_index_scatter():
    IC.check_index_scatter(DEV)


@pytest.mark.parametrize("rnn", ["LSTM", "GRU"])
def test_sparse_forward_cell(rnn):
    IC.check_sparse_forward_cell(rnn, DEV)


@pytest.mark.parametrize("rnn,hidden", [("LSTM", 250), ("GRU", 250), ("LSTM", 256)])
def test_shipped_width_against_oracle(rnn, hidden):
    case = dict(name="w", kind="hier", rnn=rnn, hidden=hidden, depthT=1, depthG=5, batch=8, data_seed=31,
                weight_seed=32, vocab=(60, 80), mean_motifs=6.0)
    IC.check_against_oracle(case, DEV, max_steps=8)


def test_motif_variant_against_oracle():
    case = dict(name="m", kind="motif", rnn="LSTM", hidden=250, depthT=1, depthG=0, batch=8, data_seed=33,
                weight_seed=34, vocab=(60, 80), mean_motifs=6.0)
    IC.check_against_oracle(case, DEV)


def test_cpu_tables_are_rejected():
    import ggpm_b200
    cell = ggpm_b200.LSTM(5, 8, 2).to(DEV)
    h = (torch.zeros(6, 8), torch.zeros(6, 8))
    with pytest.raises(RuntimeError, match="CUDA devices only"):
        cell.sparse_forward(h, torch.zeros(2, 5), torch.tensor([1, 2]), torch.zeros(2, 3, dtype=torch.long))


def test_packed_weights_are_reused_until_parameters_change():
    IC.check_packed_weight_reuse(DEV)


@pytest.mark.parametrize("M,K,width,pad_to", [(1, 1, 5, 8), (37, 4, 250, 256), (3000, 7, 250, 256), (513, 6, 64, 64)])
def test_row_kernels_against_torch(M, K, width, pad_to):
    """rows_gather / rows_scatter / rows_scatter_add on their own: ~70 % zero slots (the shared-memory row-0 path of the
    scatter-add), grids larger than one wave, unaligned table widths, zero fill up to the padded width"""
    from ggpm_b200 import functional as Fn
    gen = torch.Generator().manual_seed(M * 131 + K)
    R = max(2 * M, 16)
    table = torch.randn(R, width, generator=gen).to(DEV)
    idx = torch.randint(1, R, (M * K,), generator=gen)
    idx[torch.rand(M * K, generator=gen) < 0.7] = 0
    idx_d = idx.to(torch.int32).to(DEV)
    pairs, total = Fn.rows_gather(table, idx_d, M, K, pad_to=pad_to)
    ref = table[idx.to(DEV)]
    assert torch.equal(pairs[:, :width], ref) and float(pairs[:, width:].abs().sum()) == 0.0
    ref_sum = ref.double().reshape(M, K, width).sum(1)
    assert float((total[:, :width].double() - ref_sum).abs().max()) <= 1e-5 and float(total[:, width:].abs().sum()) == 0.0

    src = torch.randn(M * K, pad_to, generator=gen).to(DEV)
    bc = torch.randn(M, pad_to, generator=gen).to(DEV)
    acc = table.clone()
    Fn.rows_scatter_add_(acc, idx_d, M, K, src=src, bcast=bc)
    want = table.double().index_add(0, idx.to(DEV), (src[:, :width] + bc.repeat_interleave(K, 0)[:, :width]).double())
    err = float((acc.double() - want).abs().max()) / float(want.abs().max())
    assert err <= 2e-6, err

    sub = torch.randperm(R, generator=gen)[:M].to(torch.int32).to(DEV)
    rows = torch.randn(M, pad_to, generator=gen).to(DEV)
    out = Fn.rows_scatter_(table.clone(), sub, rows)
    want = table.clone()
    want[sub.long()] = rows[:, :width]
    assert torch.equal(out, want)
    assert float(Fn.rows_scatter_(table.clone(), sub, None)[sub.long()].abs().sum()) == 0.0


def test_readout_weights_are_repadded_after_a_parameter_update():
    IC.check_readout_linear_cache_invalidation(DEV)
