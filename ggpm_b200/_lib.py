"""ctypes binding of libggpm_b200.so (the C ABI declared in include/ggpm_b200.h).

There is NO CPU path: if the shared library is missing, or a tensor is not on a CUDA device, the
calls below raise.  Build the library in-tree with ``python -m ggpm_b200.build``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p

import torch

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libggpm_b200.so")

ACT_NONE, ACT_RELU, ACT_TANH, ACT_SIGMOID = 0, 1, 2, 3
ENGINE_AUTO, ENGINE_SIMT_FP32, ENGINE_TCGEN05_3XTF32, ENGINE_TCGEN05_F16X3 = 0, 1, 2, 3

P = c_void_p
I = c_int
# name -> (restype, argtypes); mirrors include/ggpm_b200.h one to one (checked by tests/test_cabi.py)
SIGNATURES = {
    "ggpm_last_error": (c_char_p, []),
    "ggpm_abi_version": (I, []),
    "ggpm_set_gemm_engine": (I, [I]),
    "ggpm_get_gemm_engine": (I, []),
    "ggpm_launch_count": (c_longlong, []),
    "ggpm_prof_enable": (I, [I]),
    "ggpm_prof_collect": (I, [P, P, P, I]),
    "ggpm_csr_from_padded": (I, [P, I, I, I, I, P, P, P, P]),
    "ggpm_check_ids": (I, [P, I, c_longlong, c_longlong, c_longlong, c_longlong, P, I, P]),
    "ggpm_csr_transpose": (I, [P, P, I, I, P, P, P, P, P]),
    "ggpm_gather_sum": (I, [P, I, P, P, P, I, I, I, I, P]),
    "ggpm_embedding_bwd": (I, [P, I, P, I, c_longlong, I, I, I, P, I, P]),
    "ggpm_gemm": (I, [P, c_longlong, c_longlong, P, c_longlong, c_longlong, P, I, I, I, I, c_float, P, I, I, I, P]),
    "ggpm_gemm_fused": (I, [P, c_longlong, c_longlong, P, c_longlong, c_longlong, P, I, I, I, I, c_float, P, I, I, I,
                            P, P, P, P, I, I, P, c_longlong, I, P, P, P]),
    "ggpm_gemm_f16x3_planes": (I, [P, P, P, c_longlong, P, P, P, c_longlong, P, I, I, I, I, c_float, I, P]),
    "ggpm_tf32_split": (I, [P, c_longlong, P, P, P]),
    "ggpm_f16_prepare_weights": (I, [P, I, I, I, P, P, P, P]),
    "ggpm_gemm_f16x3": (I, [P, c_longlong, P, P, P, c_longlong, P, I, I, I, I, c_float, P, I, I, P, P, I, I, P, c_longlong,
                            P, I, c_float, P, I, P]),
    "ggpm_lstm_pack": (I, [P, P, I, I, I, I, I, I, I, P, P, P, P]),
    "ggpm_lstm_unpack": (I, [P, P, I, I, I, I, I, I, P, P, P, P]),
    "ggpm_split_pad": (I, [P, I, I, I, I, I, P, P, P, P, P, P, P]),
    "ggpm_split_unpad": (I, [P, I, I, I, P, P, P, P, P, P, P]),
    "ggpm_act_bwd": (I, [P, I, P, I, P, I, I, I, I, I, P]),
    "ggpm_bias_act": (I, [P, I, I, I, P, I, I, P]),
    "ggpm_colsum": (I, [P, I, I, I, P, I, P]),
    "ggpm_table_wgrad": (I, [P, I, I, I, P, P, P, I, I, P]),
    "ggpm_lstm_plan": (I, [I, I]),
    "ggpm_lstm_fwd": (I, [I, I, I, I, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, I, I, P]),
    "ggpm_lstm_bwd": (I, [I, I, I, I] + [P] * 28 + [I, P, P, c_longlong, P]),
    "ggpm_gru_fwd": (I, [I, I, I, I] + [P] * 13 + [I, P]),
    "ggpm_gru_bwd": (I, [I, I, I, I] + [P] * 24),
    "ggpm_rsample_fwd": (I, [P, I, I, P, I, I, I, P, I, P, P, P]),
    "ggpm_rsample_fwd_noalloc": (I, [P, I, I, P, I, I, I, P, I, P, P]),
    "ggpm_rsample_bwd": (I, [P, I, I, P, I, I, I, P, I, P, P, I, I, P]),
    "ggpm_mlp_heads_fwd": (I, [P, I, I, I, I, I, I, P]),
    "ggpm_mlp_heads_bwd": (I, [P, P, I, I, I, I, I, I, I, P]),
    "ggpm_rows_gather": (I, [P, I, P, I, I, I, I, P, I, P, I, P]),
    "ggpm_rows_scatter": (I, [P, I, P, I, I, P, I, P]),
    "ggpm_rows_scatter_add": (I, [P, I, P, I, I, I, P, I, P, I, P]),
    "ggpm_sparse_lstm_cell_fwd": (I, [I, I, I, P, I, P, P, P, P, P, P, P]),
    "ggpm_sparse_lstm_cell_bwd": (I, [I, I, I, P, P, P, P, P, P, P, P, P, P]),
    "ggpm_sparse_gru_gate_fwd": (I, [I, I, I, P, I, P, P, P, P, P]),
    "ggpm_sparse_gru_gate_bwd": (I, [I, I, I, P, P, P, P, P, P, I, P]),
    "ggpm_sparse_gru_out_fwd": (I, [I, I, P, I, P, P, P, P]),
    "ggpm_sparse_gru_out_bwd": (I, [I, I, P, P, P, P, P, P, I, P]),
    "ggpm_sparse_lstm_fwd": (I, [I, I, I, I, I, P, P, I] + [P] * 13),
    "ggpm_sparse_lstm_bwd": (I, [I, I, I, I, I, P, P, I] + [P] * 18),
    "ggpm_sparse_gru_fwd": (I, [I, I, I, I, I, P, I] + [P] * 14),
    "ggpm_sparse_gru_bwd": (I, [I, I, I, I, I, P, I] + [P] * 21),
    "ggpm_tensorize_graph": (I, [I, P, P, I, P, P, P]),
    "ggpm_tensorize_fetch": (I, [P] * 9),
    "ggpm_tensorize_cgraph": (I, [I, I, P, P, P, P, P, P, P]),
}

_LIB = None
_TEST_BACKEND = None      # tests/ may install an emulator of the C ABI; the product never does


def _load():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "ggpm_b200: %s is missing — build it with `python -m ggpm_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)        # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB


def backend():
    return _TEST_BACKEND if _TEST_BACKEND is not None else _load()


def _install_test_backend(obj):
    """TEST SEAM ONLY (tests/_emu_lib.py): route the C-ABI calls to a host emulator.  Refused unless the process was
    started with GGPM_ALLOW_TEST_BACKEND=1 (tests/conftest.py sets it; bench.py aborts when it is set), so no product
    run can end up on a CPU route by accident."""
    global _TEST_BACKEND
    if obj is not None and os.environ.get("GGPM_ALLOW_TEST_BACKEND") != "1":
        raise RuntimeError("ggpm_b200: the C-ABI emulator seam is for tests only (set GGPM_ALLOW_TEST_BACKEND=1)")
    _TEST_BACKEND = obj


def call(name, *args):
    """Invoke a C-ABI entry point; non-zero status -> RuntimeError with ggpm_last_error()."""
    lib = backend()
    status = getattr(lib, name)(*args)
    if status != 0:
        msg = lib.ggpm_last_error()
        if isinstance(msg, bytes):
            msg = msg.decode(errors="replace")
        raise RuntimeError("%s failed (status %d): %s" % (name, status, msg))


def call_host(name, *args):
    """Invoke a host-only entry point of the real library (the index builder): never routed to a test backend."""
    lib = _load()
    status = getattr(lib, name)(*args)
    if status != 0:
        raise RuntimeError("%s failed (status %d): %s" % (name, status, lib.ggpm_last_error().decode(errors="replace")))


def require_device(t: torch.Tensor, what: str = "tensor"):
    if _TEST_BACKEND is None and not t.is_cuda:
        raise RuntimeError("ggpm_b200 runs on CUDA devices only (no CPU fallback): %s is on %s" % (what, t.device))


def fptr(t: torch.Tensor, what: str = "tensor"):
    """device pointer of an fp32 tensor whose last dimension is contiguous"""
    require_device(t, what)
    if t.dtype != torch.float32:
        raise ValueError("%s must be float32, got %s" % (what, t.dtype))
    if t.dim() >= 1 and t.shape[-1] > 1 and t.stride(-1) != 1:
        raise ValueError("%s must be contiguous along its last dimension" % what)
    return t.data_ptr()


def iptr(t: torch.Tensor, what: str = "index"):
    require_device(t, what)
    if t.dtype != torch.int32 or not t.is_contiguous():
        raise ValueError("%s must be a contiguous int32 tensor" % what)
    return t.data_ptr()


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device=None):
    """cudaStream_t of torch's current stream on `device` (the library launches on it)"""
    if _TEST_BACKEND is not None:
        return None
    if _RAW_STREAM is not None:          # same value as current_stream(device).cuda_stream without building a Stream object
        if isinstance(device, torch.device) and device.index is not None:
            return _RAW_STREAM(device.index)
        return _RAW_STREAM(torch.cuda.current_device() if device is None or getattr(device, "index", None) is None
                           else device.index)
    return torch.cuda.current_stream(device).cuda_stream


def launch_count() -> int:
    return int(backend().ggpm_launch_count())


def set_gemm_engine(engine: int):
    call("ggpm_set_gemm_engine", engine)
