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
