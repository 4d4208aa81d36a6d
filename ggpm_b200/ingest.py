"""Batch ingest for the encoder path — replaces the reference's per-step host work between the pickled batch and
the encoder call:

    ggpm/nnutils.py:201-214   make_tensor / make_cuda: ndarray -> .tolist() -> torch.tensor -> .to(device) -> .long()
                              per tensor (nine pageable copies and nine widening kernels per step)
    ggpm/dataset.py:95-118    DataFolder yields the pickled (tree_tensors, graph_tensors) tuples

``PackedBatch`` is prepared ONCE per batch (at preprocess time or in a loader thread): every index array of
``(tree_tensors, graph_tensors)`` as int32, together with what the encoder derives from them on every step — the CSR
and reverse CSR of agraph / bgraph / cgraph, the per-message source / one-hot column ids, the per-node out-message
lists and the root ids — laid out in ONE pinned host buffer.  ``to_device()`` is then a single asynchronous
host-to-device copy; the tensors it returns are views of that one device buffer and carry the prepared structures,
so ``HierMPNEncoder.forward`` / ``MotifEncoder.forward`` launch no index-preparation kernel at all.  The tensors keep
the reference layout (same shapes, same values, int32 instead of int64 — torch indexing accepts both), so they can
also be handed to code that expects ``make_cuda``'s output.

The host-side CSR builders below define the same encoding as the device builders in csrc/csr.cu (every non-zero
entry in slot order; reverse lists ascending by destination row, ties in forward order) and are checked against them
bit for bit (tests/test_ingest.py, tests/test_gpu_ingest.py).
"""
from __future__ import annotations

import io
import json
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import functional as Fn

N_BOND_TYPES = 4     # len(MolGraph.BOND_LIST), ggpm/mol_graph.py:14
MAX_POS = 20         # MolGraph.MAX_POS, ggpm/mol_graph.py:15
_ALIGN = 4           # int32 elements: every segment starts on a 16-byte boundary
_MAGIC = b"GGPMPB01"


# ----------------------------------------------------------------------------------------------
# host CSR builders (numpy, int32)
# ----------------------------------------------------------------------------------------------
def csr_from_padded_host(index: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """(row_ptr [rows+1], col [nnz]) of the non-zero entries of a padded [rows, width] index array, slot order
    (create_pad_tensor layout, ggpm/nnutils.py:105-110; zeros may also sit mid-row, ggpm/decoder.py:72-83)."""
    index = np.asarray(index)
    if index.ndim == 1:
        index = index[:, None]
    mask = index != 0
    row_ptr = np.zeros(index.shape[0] + 1, dtype=np.int32)
    np.cumsum(mask.sum(axis=1), out=row_ptr[1:])
    col = index[mask].astype(np.int32)          # boolean indexing walks row-major: slot order
    return row_ptr, col


def csr_transpose_host(row_ptr: np.ndarray, col: np.ndarray, n_src: int):
    """Reverse lists: for every source row n the destination rows e that list n (ascending, duplicates kept) and the
    position of each (e, n) pair in the forward ``col`` array.  Returns (t_ptr [n_src+1], t_row [nnz], t_pos [nnz])."""
    rows = row_ptr.shape[0] - 1
    nnz = int(row_ptr[-1])
    col = col[:nnz]
    if nnz and (col.min() < 0 or col.max() >= n_src):
        raise IndexError("neighbour index out of range [0, %d)" % n_src)
    order = np.argsort(col, kind="stable").astype(np.int32)      # forward order is (e, slot) ascending: ties keep it
    row_of = np.repeat(np.arange(rows, dtype=np.int32), np.diff(row_ptr))
    t_ptr = np.zeros(n_src + 1, dtype=np.int32)
    np.cumsum(np.bincount(col, minlength=n_src), out=t_ptr[1:])
    return t_ptr, row_of[order], order


# ----------------------------------------------------------------------------------------------
# the packed batch
# ----------------------------------------------------------------------------------------------
class _Layout:
    """name -> (offset, shape) inside one flat int32 buffer"""

    def __init__(self):
        self.entries: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        self.size = 0
        self._arrays: List[Tuple[int, np.ndarray]] = []

    def add(self, name: str, arr: np.ndarray):
        arr = np.ascontiguousarray(arr, dtype=np.int32)
        self.entries[name] = (self.size, tuple(int(s) for s in arr.shape))
        self._arrays.append((self.size, arr))
        self.size += (arr.size + _ALIGN - 1) // _ALIGN * _ALIGN

    def fill(self, flat: np.ndarray):
        for off, arr in self._arrays:
            flat[off:off + arr.size] = arr.reshape(-1)
        self._arrays = []


def _add_csr(lay: _Layout, name: str, index: np.ndarray, n_src: int, transpose: bool):
    row_ptr, col = csr_from_padded_host(index)
    if col.size and (col.min() < 0 or col.max() >= n_src):
        raise IndexError("%s: neighbour index out of range [0, %d)" % (name, n_src))
    lay.add(name + ".ptr", row_ptr)
    lay.add(name + ".col", col if col.size else np.zeros(1, np.int32))
    if transpose:
        t_ptr, t_row, t_pos = csr_transpose_host(row_ptr, col, n_src)
        lay.add(name + ".t_ptr", t_ptr)
        lay.add(name + ".t_row", t_row if t_row.size else np.zeros(1, np.int32))
        lay.add(name + ".t_pos", t_pos if t_pos.size else np.zeros(1, np.int32))


class ScopeList(list):
    """`scope` as the reference passes it (list of (start, length)), carrying the device copy of the root ids."""
    roots_dev: Optional[torch.Tensor] = None


class PackedBatch:
    """One batch ready for the device.  Build with ``PackedBatch.from_tensors(tree_tensors, graph_tensors)`` from what
    ``MolGraph.tensorize`` returns (ggpm/mol_graph.py:199-236; numpy arrays or CPU tensors; ``graph_tensors`` may be
    None for the motif-only encoder), ship with ``to_device()``, store with ``save()`` / ``load()``."""

    def __init__(self, host: torch.Tensor, entries, tree_scope, graph_scope):
        self.host = host                                  # flat int32, pinned when a CUDA device is present
        self.entries = entries                            # name -> (offset, shape)
        self.tree_scope = [tuple(int(v) for v in s) for s in tree_scope]
        self.graph_scope = None if graph_scope is None else [tuple(int(v) for v in s) for s in graph_scope]

    # -- construction ------------------------------------------------------------------------------------
    @classmethod
    def from_tensors(cls, tree_tensors, graph_tensors=None, pin: Optional[bool] = None) -> "PackedBatch":
        as_np = lambda a: a.cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
        t_fnode, t_fmess, t_agraph, t_bgraph, t_cgraph = [as_np(a) for a in tree_tensors[:5]]
        tree_scope = tree_tensors[5]
        lay = _Layout()
        n_tm = t_fmess.shape[0]

        def in_range(col, hi, what):
            # the ids below index tables inside the cell kernels; the reference's index_select raises on the same
            # violations (ggpm/encoder.py:103-106, 114-116, 121-125).  Neighbour tables are checked by _add_csr.
            col = np.asarray(col)
            if col.size and (int(col.min()) < 0 or int(col.max()) >= hi):
                raise IndexError("malformed batch: %s outside [0, %d)" % (what, hi))
        in_range(t_fmess[:, 2], MAX_POS, "tree message position (fmess[:, 2])")
        in_range([st for st, _ in tree_scope], t_fnode.shape[0], "scope start")
        lay.add("tree.fnode", t_fnode)
        lay.add("tree.fmess", t_fmess)
        lay.add("tree.agraph", t_agraph)
        lay.add("tree.bgraph", t_bgraph)
        lay.add("tree.cgraph", t_cgraph)
        _add_csr(lay, "tree.agraph", t_agraph, n_tm, True)
        _add_csr(lay, "tree.bgraph", t_bgraph, n_tm, True)
        lay.add("tree.src", t_fmess[:, 0])
        lay.add("tree.k1", t_fmess[:, 2])
        _add_csr(lay, "tree.by_src", t_fmess[:, 0], t_fnode.shape[0], True)
        lay.add("tree.roots", np.asarray([st for st, _ in tree_scope], dtype=np.int32))
        graph_scope = None
        if graph_tensors is not None:
            g_fnode, g_fmess, g_agraph, g_bgraph = [as_np(a) for a in graph_tensors[:4]]
            graph_scope = graph_tensors[4]
            n_gm = g_fmess.shape[0]
            in_range(g_fmess[:, 2], N_BOND_TYPES, "bond type (fmess[:, 2])")
            in_range(g_fmess[:, 3], MAX_POS, "bond position (fmess[:, 3])")
            lay.add("graph.fnode", g_fnode)
            lay.add("graph.fmess", g_fmess)
            lay.add("graph.agraph", g_agraph)
            lay.add("graph.bgraph", g_bgraph)
            _add_csr(lay, "graph.agraph", g_agraph, n_gm, True)
            _add_csr(lay, "graph.bgraph", g_bgraph, n_gm, True)
            _add_csr(lay, "tree.cgraph", t_cgraph, g_fnode.shape[0], True)
            lay.add("graph.src", g_fmess[:, 0])
            lay.add("graph.k1", g_fmess[:, 2])
            lay.add("graph.k2", g_fmess[:, 3] + N_BOND_TYPES)
            _add_csr(lay, "graph.by_src", g_fmess[:, 0], g_fnode.shape[0], True)
        host = torch.empty(max(lay.size, 1), dtype=torch.int32)
        if pin is None:
            pin = torch.cuda.is_available()
        if pin:
            host = host.pin_memory()
        host.zero_()
        lay.fill(host.numpy())
        return cls(host, lay.entries, tree_scope, graph_scope)

    @property
    def nbytes(self) -> int:
        return self.host.numel() * 4

    # -- wire format: magic, header length, JSON header, raw little-endian int32 payload --------------------
    def save(self, f):
        own = isinstance(f, (str, bytes))
        fh = open(f, "wb") if own else f
        header = json.dumps({"entries": {k: [v[0], list(v[1])] for k, v in self.entries.items()},
                             "tree_scope": self.tree_scope, "graph_scope": self.graph_scope,
                             "size": int(self.host.numel())}).encode()
        fh.write(_MAGIC)
        fh.write(np.asarray([len(header)], dtype="<u8").tobytes())
        fh.write(header)
        fh.write(self.host.numpy().astype("<i4", copy=False).tobytes())
        if own:
            fh.close()

    @classmethod
    def load(cls, f, pin: Optional[bool] = None) -> "PackedBatch":
        own = isinstance(f, (str, bytes))
        fh = open(f, "rb") if own else f
        if fh.read(len(_MAGIC)) != _MAGIC:
            raise ValueError("not a ggpm_b200 packed batch")
        hlen = int(np.frombuffer(fh.read(8), dtype="<u8")[0])
        header = json.loads(fh.read(hlen).decode())
        size = int(header["size"])
        host = torch.empty(size, dtype=torch.int32)
        if pin is None:
            pin = torch.cuda.is_available()
        if pin:
            host = host.pin_memory()
        payload = fh.read(4 * size)
        if len(payload) != 4 * size:
            raise ValueError("truncated packed batch")
        host.numpy()[:] = np.frombuffer(payload, dtype="<i4")
        if own:
            fh.close()
        entries = {k: (int(v[0]), tuple(int(s) for s in v[1])) for k, v in header["entries"].items()}
        return cls(host, entries, header["tree_scope"], header["graph_scope"])

    # -- to the device ------------------------------------------------------------------------------------
    def _view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        off, shape = self.entries[name]
        n = int(np.prod(shape)) if len(shape) else 1
        return buf[off:off + n].view(shape)

    def _csr(self, buf, name: str, rows: int, n_src: int) -> Fn.CSR:
        csr = Fn.CSR(self._view(buf, name + ".ptr"), self._view(buf, name + ".col"), rows, n_src)
        t = Fn.CSR(self._view(buf, name + ".t_ptr"), self._view(buf, name + ".t_row"), n_src, rows)
        t.pos = self._view(buf, name + ".t_pos")
        csr._t, t._t = t, csr
        return csr

    def to_device(self, device=None, non_blocking: bool = True):
        """ONE host-to-device copy; returns (tree_tensors, graph_tensors) in the reference layout (int32 views of the
        device buffer) with the prepared index structures attached (graph_tensors is None for a tree-only batch)."""
        if device is None:
            from .nnutils import default_device
            device = default_device()
        buf = self.host.to(device, non_blocking=non_blocking)
        return self.views(buf)

    def views(self, buf: torch.Tensor):
        v = lambda name: self._view(buf, name)
        t_fnode, t_fmess, t_agraph, t_bgraph, t_cgraph = (v("tree." + k) for k in ("fnode", "fmess", "agraph", "bgraph", "cgraph"))
        n_tn, n_tm = t_fnode.shape[0], t_fmess.shape[0]
        t_agraph._ggpm_csr = self._csr(buf, "tree.agraph", n_tn, n_tm)
        t_bgraph._ggpm_csr = self._csr(buf, "tree.bgraph", n_tm, n_tm)
        t_fmess._ggpm_cols = (v("tree.src"), v("tree.k1"), None)
        t_fmess._ggpm_by_src = self._csr(buf, "tree.by_src", n_tm, n_tn).T
        tree_scope = ScopeList(self.tree_scope)
        tree_scope.roots_dev = v("tree.roots")
        tree_tensors = [t_fnode, t_fmess, t_agraph, t_bgraph, t_cgraph, tree_scope]
        if self.graph_scope is None:
            return tree_tensors, None
        g_fnode, g_fmess, g_agraph, g_bgraph = (v("graph." + k) for k in ("fnode", "fmess", "agraph", "bgraph"))
        n_gn, n_gm = g_fnode.shape[0], g_fmess.shape[0]
        g_agraph._ggpm_csr = self._csr(buf, "graph.agraph", n_gn, n_gm)
        g_bgraph._ggpm_csr = self._csr(buf, "graph.bgraph", n_gm, n_gm)
        g_fmess._ggpm_cols = (v("graph.src"), v("graph.k1"), v("graph.k2"))
        g_fmess._ggpm_by_src = self._csr(buf, "graph.by_src", n_gm, n_gn).T
        t_cgraph._ggpm_csr = self._csr(buf, "tree.cgraph", n_tn, n_gn)
        graph_tensors = [g_fnode, g_fmess, g_agraph, g_bgraph, ScopeList(self.graph_scope)]
        return tree_tensors, graph_tensors


def pack(tensors, pin: Optional[bool] = None) -> PackedBatch:
    """(tree_tensors, graph_tensors) -> PackedBatch (the argument make_cuda takes, ggpm/nnutils.py:210)"""
    tree_tensors, graph_tensors = tensors
    return PackedBatch.from_tensors(tree_tensors, graph_tensors, pin=pin)


def make_cuda_packed(tensors, device=None):
    """Drop-in for make_cuda (ggpm/nnutils.py:210-214) on a batch that was not packed ahead of time: pack, one copy."""
    return pack(tensors).to_device(device)


class DeviceLoader:
    """Keeps ``depth`` packed batches in flight: the host-to-device copy of batch i+1 runs on a side stream while batch
    i is being encoded (what the DataLoader + make_cuda pair of the reference cannot do: its copies are pageable and
    its index tensors are widened on the compute stream, ggpm/nnutils.py:201-214).  Iterating yields
    ``(tree_tensors, graph_tensors)`` exactly like ``PackedBatch.to_device()``; the compute stream waits only for the
    copy of the batch it is about to use."""

    def __init__(self, batches, device=None, depth: int = 2):
        if device is None:
            from .nnutils import default_device
            device = default_device()
        self.device = torch.device(device)
        self.depth = max(1, int(depth))
        self._it = iter(batches)
        self._queue = []
        self._copy_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def _enqueue(self) -> bool:
        try:
            pb = next(self._it)
        except StopIteration:
            return False
        if self._copy_stream is None:
            self._queue.append((pb, pb.host.to(self.device), None))
            return True
        with torch.cuda.stream(self._copy_stream):
            buf = pb.host.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._queue.append((pb, buf, ev))
        return True

    def __iter__(self):
        return self

    def __next__(self):
        while len(self._queue) < self.depth and self._enqueue():
            pass
        if not self._queue:
            raise StopIteration
        pb, buf, ev = self._queue.pop(0)
        if ev is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            buf.record_stream(cur)          # the buffer was allocated on the copy stream and is consumed on this one
        self._enqueue()                     # keep the pipeline full before the caller launches its step
        return pb.views(buf)
