"""CPU: the product's Python host side (index prep, weight packing, autograd wiring, pointer and stride
arithmetic of every C-ABI call) driven end to end against a host EMULATOR of the C ABI
(tests/_emu_lib.py), and compared with the golden vectors produced by the unmodified reference.
This does not test the CUDA kernels (the -m gpu tests do); it pins everything around them."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import golden_names                                     # noqa: E402
import _checks as C                                                    # noqa: E402
import _emu_lib                                                        # noqa: E402
from ggpm_b200 import _lib, nnutils                                    # noqa: E402

DEV = "cpu"


@pytest.fixture()
def emu():
    lib = _emu_lib.EmuLib()
    _lib._install_test_backend(lib)
    nnutils.set_default_device("cpu")
    yield lib
    _lib._install_test_backend(None)
    nnutils.set_default_device(None)


SMALL = [n for n in golden_names() if n != "hier_lstm_cfg1"]


@pytest.mark.parametrize("name", SMALL)
def test_forward_backward_matches_reference_golden(emu, name):
    C.check_forward_backward_golden(name, DEV)


@pytest.mark.parametrize("name", ["hier_lstm_small", "hier_gru_small", "motif_lstm_small"])
def test_packed_batch_path_matches_reference_golden(emu, name):
    C.check_forward_backward_golden(name, DEV, packed=True)
    # the packed path prepares every index structure on the host: no CSR kernels are launched
    assert not [c for c in emu.calls if "csr" in c], sorted(set(emu.calls))


def test_root_only_loss_leaves_tree_readout_without_grad(emu):
    C.check_root_only_loss(DEV)


def test_no_grad_and_eval_mode(emu):
    C.check_no_grad_eval(DEV)
    assert "lstm_bwd" not in emu.calls and "gru_bwd" not in emu.calls


def test_train_mode_dropout_is_applied(emu):
    C.check_train_dropout(DEV)


@pytest.mark.parametrize("name", ["hier_lstm_small", "hier_gru_small", "motif_lstm_small"])
def test_train_mode_parity_with_fixed_dropout_masks(emu, name):
    C.check_fixed_mask_dropout_parity(name, DEV, p=0.1)


def test_malformed_batches_raise_index_error(emu):
    C.check_malformed_batches_raise(DEV)


def test_inputs_are_not_mutated(emu):
    C.check_inputs_not_mutated(DEV)


@pytest.mark.parametrize("rnn", ["LSTM", "GRU"])
def test_mpn_encoder_reference_signature(emu, rnn):
    C.check_mpn_reference_signature(rnn, DEV)


def test_lstm_cell_state_gradient(emu):
    C.check_lstm_cell_state_gradient(DEV)


def test_index_select_nd(emu):
    C.check_index_select_nd(DEV)


def test_cpu_tensors_rejected_without_backend():
    _lib._install_test_backend(None)
    with pytest.raises(RuntimeError, match="CUDA devices only"):
        _lib.require_device(torch.zeros(3))


def test_modules_refuse_to_build_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import ggpm_b200
    from oracle.ref_loader import FakePairVocab, IdAtomVocab
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ggpm_b200.HierMPNEncoder(FakePairVocab(10, 10), IdAtomVocab(), "LSTM", 8, 8, 2, 2, 0.0)
