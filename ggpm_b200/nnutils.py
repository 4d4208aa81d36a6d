"""Tensor utilities of the encoder path — host-side mirror of ggpm/nnutils.py for the functions the
encoder uses: index_select_ND (:65-70), create_pad_tensor (:105-110), index_scatter (:124-128),
make_tensor / make_cuda (:201-214).  Device work goes through libggpm_b200.so; there is no CPU implementation."""
from __future__ import annotations

import numpy as np
import torch

from . import functional as Fn

_DEVICE = None


def default_device() -> torch.device:
    """The reference fixes a global device at import time (ggpm/nnutils.py:9-10: cuda:0 if available).
    One process per GPU: cuda:<current device> of this process."""
    if _DEVICE is not None:
        return _DEVICE
    if not torch.cuda.is_available():
        raise RuntimeError("ggpm_b200 needs a CUDA device (the encoder path has no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def set_default_device(device):
    global _DEVICE
    _DEVICE = None if device is None else torch.device(device)


def to_cuda(tensor):
    return tensor.to(default_device())


class _PinnedRing:
    """Small ring of pinned int64 staging buffers for host lists that must reach the device every step (the root ids
    derived from `scope`).  A pageable torch.tensor(list, device=...) copy makes the host wait for the stream, which
    serialises host and device once per forward; a pinned non_blocking copy does not.  Each slot carries an event so
    that it is only rewritten after its previous copy has executed."""

    def __init__(self, slots: int = 8):
        self.slots, self.buf, self.ev, self.i = slots, [None] * slots, [None] * slots, 0

    def to_device(self, values, device) -> torch.Tensor:
        n = len(values)
        k = self.i
        self.i = (self.i + 1) % self.slots
        if self.ev[k] is not None:
            self.ev[k].synchronize()
        if self.buf[k] is None or self.buf[k].numel() < n:
            self.buf[k] = torch.empty(max(256, 2 * n), dtype=torch.int64).pin_memory()
        host = self.buf[k][:n]
        host.numpy()[:] = np.asarray(values, dtype=np.int64)
        out = host.to(device, non_blocking=True)
        if self.ev[k] is None:
            self.ev[k] = torch.cuda.Event()
        self.ev[k].record(torch.cuda.current_stream(device))
        return out


_RING = None


def ints_to_device(values, device) -> torch.Tensor:
    """Python list of ints -> int64 tensor on `device` without a host/stream synchronisation (CUDA devices);
    plain torch.tensor elsewhere (the host-side tests)."""
    global _RING
    device = torch.device(device)
    if device.type != "cuda":
        return torch.tensor(list(values), dtype=torch.long, device=device)
    if _RING is None:
        _RING = _PinnedRing()
    return _RING.to_device(values, device)


def index_select_ND(source: torch.Tensor, dim: int, index: torch.Tensor) -> torch.Tensor:
    """source [R, H], index [M, K] -> [M, K, H]; same semantics as ggpm/nnutils.py:65-70 (padding slots
    read row `index` points at, normally the zero row 0).  Materialises the result — the encoder itself
    uses index_select_sum below, which never does."""
    if dim != 0 or source.dim() != 2:
        raise ValueError("index_select_ND supports dim=0 on a 2-D source")
    flat = index.reshape(-1).to(torch.int32)
    n = flat.numel()
    row_ptr = torch.arange(n + 1, dtype=torch.int32, device=source.device)
    csr = Fn.CSR(row_ptr, flat.contiguous(), n, source.shape[0])
    return Fn.gather_sum(source, csr).reshape(tuple(index.shape) + (source.shape[1],))


def index_select_sum(source: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    """index_select_ND(source, 0, index).sum(dim=1) as one CSR gather-sum kernel."""
    return Fn.gather_sum(source, Fn.csr_from_padded(index, source.shape[0]))


def index_scatter(sub_data, all_data, index):
    """all_data with rows `index` replaced by sub_data (ggpm/nnutils.py:124-128), differentiable in both"""
    return Fn.index_scatter(sub_data, all_data, index)


def create_pad_tensor(alist):
    """list of id lists -> zero padded IntTensor with one extra trailing zero column (ggpm/nnutils.py:105-110)"""
    width = max(len(a) for a in alist) + 1
    out = np.zeros((len(alist), width), dtype=np.int32)
    for i, a in enumerate(alist):
        out[i, :len(a)] = a
    return torch.from_numpy(out)


def make_tensor(x):
    if not isinstance(x, torch.Tensor):
        x = torch.from_numpy(np.ascontiguousarray(x)) if isinstance(x, np.ndarray) else torch.tensor(x)
    return to_cuda(x)


def make_cuda(tensors):
    """(tree_tensors, graph_tensors) of numpy / tensors -> torch.long tensors on the device; the trailing
    `scope` list is passed through (ggpm/nnutils.py:210-214)."""
    tree_tensors, graph_tensors = tensors
    tree_tensors = [make_tensor(x).long() for x in tree_tensors[:-1]] + [tree_tensors[-1]]
    graph_tensors = [make_tensor(x).long() for x in graph_tensors[:-1]] + [graph_tensors[-1]]
    return tree_tensors, graph_tensors
