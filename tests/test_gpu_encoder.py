"""B200: the real CUDA library behind the drop-in modules, against the golden vectors of the unmodified
reference and the oracle on seeded inputs (fp32 path, tolerance 1e-5 relative on outputs as BASELINE.json's
north_star states, on outputs and on gradients).  Every call goes through the C ABI (ggpm_b200/_lib.py)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import golden_names                                     # noqa: E402
import _checks as C                                                    # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"
SMALL = [n for n in golden_names() if n != "hier_lstm_cfg1"]
ENGINES = [1, 2, 0]   # 1 = fp32 CUDA cores only, 2 = tcgen05 3xTF32, 0 = auto (f16x3 for the LSTM gate GEMMs, 3xTF32 elsewhere)


@pytest.fixture(params=ENGINES, ids=["simt", "tf32x3", "auto"])
def engine(request):
    from ggpm_b200 import _lib
    assert _lib._TEST_BACKEND is None
    _lib.set_gemm_engine(request.param)
    yield request.param
    _lib.set_gemm_engine(0)


@pytest.mark.parametrize("name", SMALL)
def test_forward_backward_matches_reference_golden(engine, name):
    C.check_forward_backward_golden(name, DEV)


def test_cfg1_matches_reference_golden(engine):
    C.check_cfg1(DEV)


def test_root_only_loss_leaves_tree_readout_without_grad(engine):
    C.check_root_only_loss(DEV)


def test_no_grad_and_eval_mode(engine):
    C.check_no_grad_eval(DEV)


def test_train_mode_dropout_is_applied():
    C.check_train_dropout(DEV)


@pytest.mark.parametrize("name", ["hier_lstm_small", "hier_lstm_deep", "hier_gru_small", "motif_lstm_small"])
@pytest.mark.parametrize("p", [0.1, 0.5])
def test_train_mode_parity_with_fixed_dropout_masks(name, p):
    C.check_fixed_mask_dropout_parity(name, DEV, p=p)


def test_malformed_batches_raise_index_error():
    C.check_malformed_batches_raise(DEV)


def test_inputs_are_not_mutated():
    C.check_inputs_not_mutated(DEV)


@pytest.mark.parametrize("rnn", ["LSTM", "GRU"])
def test_mpn_encoder_reference_signature(engine, rnn):
    C.check_mpn_reference_signature(rnn, DEV)


def test_lstm_cell_state_gradient(engine):
    C.check_lstm_cell_state_gradient(DEV)


def test_index_select_nd():
    C.check_index_select_nd(DEV)


def test_cpu_tensors_are_rejected():
    import ggpm_b200
    from oracle.ref_loader import FakePairVocab, IdAtomVocab
    from conftest import load_golden
    meta, z, tree, graph = load_golden("motif_lstm_small")
    model = C.build_model(meta, DEV)
    with pytest.raises(RuntimeError, match="CUDA devices only"):
        model(tree)          # index tensors left on the host


def test_launches_are_counted():
    from ggpm_b200 import _lib
    before = _lib.launch_count()
    C.check_no_grad_eval(DEV)
    assert _lib.launch_count() > before


def test_plane_fed_forward_of_large_layers_matches_golden():
    """Layers of >= 32 768 rows form S with the sgather kernel and run the forward gate GEMMs from fp16 planes; the
    threshold is read once per process, so the golden checks are repeated in a child process with the threshold at 0."""
    import subprocess
    env = dict(os.environ, GGPM_SGATHER_MIN_ROWS="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "cfg1_matches or (forward_backward_matches and auto)"], env=env, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert " passed" in r.stdout


def test_backward_with_the_side_stream_matches_golden():
    """GGPM_WGRAD_OVERLAP=1 issues the weight-gradient GEMM of every backward LSTM round on a side stream (forked /
    joined with events) so that it runs under the next round's cell kernel.  The switch is read once per process, so
    the golden checks are repeated in a child process with the overlap on."""
    import subprocess
    env = dict(os.environ, GGPM_WGRAD_OVERLAP="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "cfg1_matches or (forward_backward_matches and auto and lstm)"], env=env, capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert " passed" in r.stdout


def test_atomic_free_backward_matches_golden():
    """GGPM_BWD_PAIRS=1 replaces the vector reductions of the backward cell kernel (dP, dC of the previous state) by per-pair
    stores gathered over the reverse CSR: same goldens.  (The whole backward pass is still not bit-reproducible: the
    split-K partial sums of the weight-gradient GEMM meet in fp32 reduce-adds.)  The switch is read once per process, so
    the checks run in a child process."""
    import subprocess
    env = dict(os.environ, GGPM_BWD_PAIRS="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "cfg1_matches or (forward_backward_matches and auto and lstm)"], env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert " passed" in r.stdout


def test_gate_nonlinearities_in_the_gemm_epilogue_match_golden():
    """GGPM_FUSED_GATES=1: on the plane-fed path the forward gate GEMM applies the x-projection, sigmoid / tanh and the row-0
    mask in its tcgen05 epilogue and the cell kernels only evaluate the per-neighbour forget gate (forward and backward read
    the activated gates from the history).  With GGPM_SGATHER_MIN_ROWS=0 every golden takes that path; both switches are
    read once per process, so the checks run in a child process."""
    import subprocess
    env = dict(os.environ, GGPM_FUSED_GATES="1", GGPM_SGATHER_MIN_ROWS="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "cfg1_matches or (forward_backward_matches and auto and lstm) or (fixed_dropout and lstm)"], env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert " passed" in r.stdout
