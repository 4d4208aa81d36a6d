"""Autograd functions over the C ABI (libggpm_b200.so): CSR index prep, neighbour gather-sum, fused
linear layers and the graph LSTM / GRU message passing with hand-written backward.

PyTorch is used for device memory (caching allocator), the stream and the autograd graph between
these nodes; every FLOP and every gathered byte of the hot path goes through the CUDA library.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import ACT_NONE, ACT_RELU, ACT_TANH, call, fptr, iptr, stream_ptr


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def padded_width(hidden: int) -> int:
    """internal row width of hidden-state tables: multiple of 32 floats (16-byte rows, K tiles of 32)"""
    return round_up(hidden, 32)


# ----------------------------------------------------------------------------------------------
# CSR neighbour lists
# ----------------------------------------------------------------------------------------------
class CSR:
    """Non-zero entries of a padded [rows, width] index tensor (ggpm/nnutils.py:105-110) in slot order,
    plus (lazily) the reverse lists used by the backward pass."""

    def __init__(self, row_ptr, col, rows, n_src):
        self.row_ptr, self.col, self.rows, self.n_src = row_ptr, col, rows, n_src
        self._t = None

    @property
    def T(self) -> "CSR":
        if self._t is None:
            dev = self.row_ptr.device
            cap = max(1, self.col.numel())
            t_ptr = torch.empty(self.n_src + 1, dtype=torch.int32, device=dev)
            t_row = torch.empty(cap, dtype=torch.int32, device=dev)
            t_pos = torch.empty(cap, dtype=torch.int32, device=dev)
            cursor = torch.empty(self.n_src, dtype=torch.int32, device=dev)
            call("ggpm_csr_transpose", iptr(self.row_ptr), iptr(self.col), self.rows, self.n_src,
                 iptr(t_ptr), iptr(t_row), iptr(t_pos), iptr(cursor), stream_ptr(dev))
            t = CSR(t_ptr, t_row, self.n_src, self.rows)
            t.pos = t_pos
            t._t = self
            self._t = t
        return self._t


class IndexGuard:
    """Range validation of the index tensors of one batch on the make_cuda path (tensors that came through
    ingest.PackedBatch were validated on the host when they were packed).  The reference's own index_select / embedding
    calls raise on a bad label, position or node id; here the ids feed hand-written kernels, so they are checked first:
    every check is one small kernel OR-ing a bit into one device flag, and `raise_if_bad()` reads the flag back ONCE
    per forward, before any kernel that indexes with these ids is launched.  GGPM_SKIP_INDEX_CHECK=1 skips the checks
    (and the read-back) for callers that validate their batches elsewhere."""

    _skip = None

    def __init__(self, device):
        import os
        if IndexGuard._skip is None:
            IndexGuard._skip = os.environ.get("GGPM_SKIP_INDEX_CHECK", "0") == "1"
        self.device = device
        self.flag = None
        self.names: List[str] = []

    def _bit(self, what: str) -> int:
        if self.flag is None:
            self.flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        if what not in self.names:
            self.names.append(what)
        return 1 << min(self.names.index(what) + 1, 30)          # bit 0: neighbour entries (csr_from_padded)

    def check(self, ids: torch.Tensor, lo: int, hi: int, what: str):
        """ids: any int32 / int64 view with a single stride (1-D, e.g. a column of fmess)"""
        if IndexGuard._skip or ids.numel() == 0:
            return
        _lib.require_device(ids, what)
        if ids.dim() != 1 or ids.dtype not in (torch.int32, torch.int64):
            raise ValueError("%s: expected a 1-D int32 / int64 id tensor" % what)
        call("ggpm_check_ids", ids.data_ptr(), int(ids.dtype == torch.int64), ids.numel(), max(ids.stride(0), 1), lo, hi,
             iptr(self._flag()), self._bit("%s outside [%d, %d)" % (what, lo, hi)), stream_ptr(ids.device))

    def _flag(self):
        if self.flag is None:
            self.flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        return self.flag

    def csr_flag(self):
        return None if IndexGuard._skip else self._flag()

    def raise_if_bad(self):
        if self.flag is None or IndexGuard._skip:
            return
        bits = int(self.flag.item())
        if bits:
            bad = ["neighbour index outside its table"] if bits & 1 else []
            bad += [n for i, n in enumerate(self.names) if bits & (1 << min(i + 1, 30))]
            raise IndexError("malformed batch: " + "; ".join(bad))


def csr_from_padded(index: torch.Tensor, n_src: int, check: bool = False, guard: Optional[IndexGuard] = None) -> CSR:
    """index: [rows, width] int64 (what make_cuda yields, ggpm/nnutils.py:210-214) or int32, on the GPU.
    Tensors that came out of ingest.PackedBatch.to_device() carry their CSR (built on the host, same encoding).
    `guard`: entries outside [0, n_src) set bit 0 of the guard's flag (read back once per forward)."""
    pre = getattr(index, "_ggpm_csr", None)
    if pre is not None and pre.n_src == n_src and pre.row_ptr.device == index.device:
        return pre
    _lib.require_device(index, "index tensor")
    if index.dtype not in (torch.int64, torch.int32):
        raise ValueError("index tensors must be int64 or int32, got %s" % index.dtype)
    if index.dim() == 1:
        index = index.unsqueeze(1)
    if index.dim() != 2:
        raise ValueError("index tensor must be 2-D [rows, width]")
    index = index.contiguous()
    rows, width = index.shape
    if rows < 1:
        raise ValueError("index tensor needs at least the padding row")
    dev = index.device
    row_ptr = torch.empty(rows + 1, dtype=torch.int32, device=dev)
    col = torch.empty(max(1, rows * width), dtype=torch.int32, device=dev)
    shared = None if guard is None else guard.csr_flag()
    bad = torch.zeros(1, dtype=torch.int32, device=dev) if shared is None else shared
    call("ggpm_csr_from_padded", index.data_ptr(), int(index.dtype == torch.int64), rows, width, n_src,
         iptr(row_ptr), iptr(col), iptr(bad), stream_ptr(dev))
    if check and int(bad.item()) != 0:
        raise IndexError("neighbour index out of range [0, %d)" % n_src)
    return CSR(row_ptr, col, rows, n_src)


def _row_stride(t: torch.Tensor) -> int:
    return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)


def _aligned_table(t: torch.Tensor) -> torch.Tensor:
    """fp32 2-D, unit column stride, row stride % 4 == 0, 16-byte aligned base (copy only if violated)"""
    if t.dim() != 2:
        raise ValueError("expected a 2-D table")
    ok = t.stride(1) == 1 and _row_stride(t) % 4 == 0 and t.data_ptr() % 16 == 0 and t.shape[1] % 4 == 0
    if ok:
        return t
    w = round_up(t.shape[1], 4)
    out = torch.zeros(t.shape[0], w, dtype=t.dtype, device=t.device)
    out[:, :t.shape[1]] = t
    return out


def _gather_sum_raw(table: torch.Tensor, csr: CSR, width: int) -> torch.Tensor:
    out = torch.empty(csr.rows, width, dtype=torch.float32, device=table.device)
    call("ggpm_gather_sum", fptr(table, "table"), _row_stride(table), iptr(csr.row_ptr), iptr(csr.col),
         fptr(out), width, csr.rows, width, 0, stream_ptr(table.device))
    return out


class GatherSum(torch.autograd.Function):
    """out[r] = sum_{j in csr row r} table[col[j]]  ==  index_select_ND(table, 0, index).sum(1)
    (ggpm/nnutils.py:65-70 + .sum(dim=1) at ggpm/rnn.py:26,86 and ggpm/encoder.py:32,100,136)."""

    @staticmethod
    def forward(ctx, table, csr):
        if table.shape[0] != csr.n_src:
            raise ValueError("table has %d rows, index expects %d" % (table.shape[0], csr.n_src))
        src = _aligned_table(table)
        ctx.csr, ctx.width, ctx.cols = csr, src.shape[1], table.shape[1]
        out = _gather_sum_raw(src, csr, src.shape[1])
        return out if out.shape[1] == table.shape[1] else out[:, :table.shape[1]]

    @staticmethod
    def backward(ctx, dout):
        g = _aligned_table(dout)
        if g.shape[1] != ctx.width:
            gg = torch.zeros(g.shape[0], ctx.width, dtype=g.dtype, device=g.device)
            gg[:, :g.shape[1]] = g
            g = gg
        dt = _gather_sum_raw(g, ctx.csr.T, ctx.width)
        return (dt if ctx.cols == ctx.width else dt[:, :ctx.cols]), None


def gather_sum(table: torch.Tensor, csr: CSR) -> torch.Tensor:
    return GatherSum.apply(table, csr)


class EmbeddingRows(torch.autograd.Function):
    """weight[idx] with the library's scatter-add backward (torch's dense embedding backward takes 150 us for the
    3 k motif rows of a cfg2 batch; this one 5 us).  idx: 1-D int32 / int64, any stride."""

    @staticmethod
    def forward(ctx, weight, idx):
        ctx.save_for_backward(idx)
        ctx.shape = tuple(weight.shape)
        return weight.index_select(0, idx if idx.is_contiguous() else idx.contiguous())

    @staticmethod
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        V, E = ctx.shape
        dout = dout if dout.stride(1) == 1 else dout.contiguous()
        dW = torch.empty(V, E, dtype=torch.float32, device=dout.device)
        _lib.require_device(idx, "embedding ids")
        call("ggpm_embedding_bwd", fptr(dout), _row_stride(dout), idx.data_ptr(), int(idx.dtype == torch.int64),
             idx.stride(0) if idx.numel() > 1 else 1, idx.numel(), E, V, fptr(dW), E, stream_ptr(dout.device))
        return dW, None


def embedding_rows(weight: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    if idx.dim() != 1 or idx.dtype not in (torch.int32, torch.int64):
        raise ValueError("embedding ids must be a 1-D int32 / int64 tensor")
    return EmbeddingRows.apply(weight, idx)


# ----------------------------------------------------------------------------------------------
# dense layers
# ----------------------------------------------------------------------------------------------
def _gemm(A, sam, sak, B, sbk, sbn, C, M, N, K, beta=0.0, bias=None, act=ACT_NONE, zero_row0=False, split_k=1):
    call("ggpm_gemm", fptr_any(A), sam, sak, fptr_any(B), sbk, sbn, fptr(C), _row_stride(C), M, N, K, float(beta),
         None if bias is None else fptr(bias), act, int(zero_row0), split_k, stream_ptr(C.device))


def fptr_any(t):
    _lib.require_device(t)
    if t.dtype != torch.float32:
        raise ValueError("expected float32, got %s" % t.dtype)
    return t.data_ptr()


def _kmajor(t: torch.Tensor) -> torch.Tensor:
    """2-D fp32 with unit stride along dim 1 (views of weights / padded tables are fine)"""
    if t.dim() != 2:
        raise ValueError("expected a 2-D operand")
    return t if (t.stride(1) == 1 or t.shape[1] == 1) else t.contiguous()


def gemm_nt(A, W, out=None, beta=0.0, bias=None, act=ACT_NONE, zero_row0=False, out_width=None):
    """out[M, :N] = act(A[M,K] @ W[N,K]^T + beta*out + bias)"""
    A, W = _kmajor(A), _kmajor(W)
    M, K = A.shape
    N = W.shape[0]
    if W.shape[1] != K:
        raise ValueError("gemm_nt: K mismatch %d vs %d" % (K, W.shape[1]))
    if out is None:
        width = N if out_width is None else out_width
        out = (torch.empty if width == N else torch.zeros)(M, width, dtype=torch.float32, device=A.device)
    _gemm(A, _row_stride(A), 1, W, 1, _row_stride(W), out, M, N, K, beta, bias, act, zero_row0)
    return out


def gemm_nn(A, B, out=None, beta=0.0):
    """out[M,N] = A[M,K] @ B[K,N] (+ beta*out), B row-major"""
    A, B = _kmajor(A), _kmajor(B)
    M, K = A.shape
    N = B.shape[1]
    if B.shape[0] != K:
        raise ValueError("gemm_nn: K mismatch")
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=A.device)
    _gemm(A, _row_stride(A), 1, B, _row_stride(B), 1, out, M, N, K, beta)
    return out


def gemm_tn(A, B, out=None, beta=0.0):
    """out[M,N] = A[K,M]^T @ B[K,N] (+ beta*out): weight gradients, split over K to fill the GPU"""
    A, B = _kmajor(A), _kmajor(B)
    K, M = A.shape
    N = B.shape[1]
    if B.shape[0] != K:
        raise ValueError("gemm_tn: K mismatch")
    if out is None:
        out = torch.zeros(M, N, dtype=torch.float32, device=A.device)
        beta = 1.0
    tiles = ((M + 127) // 128) * ((N + 127) // 128)
    splits = max(1, min((2 * 148 + tiles - 1) // tiles, (K + 255) // 256))
    if splits > 1 and beta == 0.0:
        out.zero_()
        beta = 1.0
    _gemm(A, 1, _row_stride(A), B, _row_stride(B), 1, out, M, N, K, beta, split_k=splits)
    return out


def gemm_fused(A, sam, sak, B, sbk, sbn, C, M, N, K, beta=0.0, a_csr: Optional[CSR] = None, b_csr: Optional[CSR] = None,
               split_dim=0, split_at=0, A2=None, lda2=0, split_k=1, tile_n=0, B_planes=None):
    """Raw tcgen05 entry (ggpm_gemm_fused): C[M,N] (+)= A(M,K) B(K,N) with CSR gather-sum operand prologues and
    the second-region switch; see include/ggpm_b200.h.  Used by the library internally for the LSTM layer; exposed
    for tests and profiling."""
    call("ggpm_gemm_fused", fptr_any(A), sam, sak, fptr_any(B), sbk, sbn, fptr(C), _row_stride(C), M, N, K, float(beta),
         None, ACT_NONE, 0, split_k,
         None if a_csr is None else iptr(a_csr.row_ptr), None if a_csr is None else iptr(a_csr.col),
         None if b_csr is None else iptr(b_csr.row_ptr), None if b_csr is None else iptr(b_csr.col),
         split_dim, split_at, None if A2 is None else fptr_any(A2), lda2, tile_n,
         None if B_planes is None else fptr(B_planes[0]), None if B_planes is None else fptr(B_planes[1]),
         stream_ptr(C.device))
    return C


def f16_prepare_weights(W: torch.Tensor, transpose: bool = False):
    """(planes, scal): fp16 hi / lo planes [2, N, K] of W * scale (of W^T when `transpose`) and the device scalars
    [max|W|, scale] (ggpm_f16_prepare_weights)"""
    W = W.contiguous()
    rows, cols = W.shape
    shape = (cols, rows) if transpose else (rows, cols)
    planes = torch.empty((2,) + shape, dtype=torch.float16, device=W.device)
    scal = torch.empty(2, dtype=torch.float32, device=W.device)
    call("ggpm_f16_prepare_weights", fptr(W), rows, cols, int(transpose), planes[0].data_ptr(), planes[1].data_ptr(),
         fptr(scal), stream_ptr(W.device))
    return planes, scal


def gemm_f16x3(A, lda, planes, scal, C, M, N, K, beta=0.0, bias=None, act=ACT_NONE, zero_row0=False,
               a_csr: Optional[CSR] = None, split_dim=0, split_at=0, A2=None, lda2=0, a_amax=None, a_amax_const=1.0,
               a_deg=None, tile_n=0):
    """Raw f16x3 entry (ggpm_gemm_f16x3): C = epilogue(A(M,K) @ B(K,N)) with B the prepared fp16 planes [2, N, K]"""
    call("ggpm_gemm_f16x3", fptr_any(A), lda, planes[0].data_ptr(), planes[1].data_ptr(), scal[1:].data_ptr(),
         planes.stride(1), fptr(C), _row_stride(C), M, N, K, float(beta), None if bias is None else fptr(bias), act,
         int(zero_row0), None if a_csr is None else iptr(a_csr.row_ptr), None if a_csr is None else iptr(a_csr.col),
         split_dim, split_at, None if A2 is None else fptr_any(A2), lda2,
         None if a_amax is None else fptr(a_amax), 0 if a_amax is None else a_amax.numel(), float(a_amax_const),
         None if a_deg is None else iptr(a_deg), tile_n, stream_ptr(C.device))
    return C


def gemm_f16x3_planes(a_planes, a_scal, b_planes, b_scal, C, M, N, K, beta=0.0, tn=False):
    """Raw entry (ggpm_gemm_f16x3_planes): both operands as prepared fp16 planes ([2, M, K] x [2, N, K], or with
    tn=True [2, K, M] x [2, K, N] for C = A^T B)"""
    call("ggpm_gemm_f16x3_planes", a_planes[0].data_ptr(), a_planes[1].data_ptr(), a_scal[1:].data_ptr(),
         a_planes.stride(1), b_planes[0].data_ptr(), b_planes[1].data_ptr(), b_scal[1:].data_ptr(), b_planes.stride(1),
         fptr(C), _row_stride(C), M, N, K, float(beta), int(tn), stream_ptr(C.device))
    return C


def tf32_split(W: torch.Tensor):
    """(hi, lo) tf32 planes of a contiguous fp32 weight matrix (ggpm_tf32_split)"""
    W = W.contiguous()
    planes = torch.empty((2,) + tuple(W.shape), dtype=torch.float32, device=W.device)
    call("ggpm_tf32_split", fptr(W), W.numel(), fptr(planes[0]), fptr(planes[1]), stream_ptr(W.device))
    return planes


class LinearAct(torch.autograd.Function):
    """Y = act(sum_i A_i @ W_i^T + b), optional zero row 0.  Replaces cat + nn.Linear + ReLU/Tanh + mask
    (ggpm/encoder.py:33-38, 101, 112, 137-138) without building the concatenation."""

    @staticmethod
    def forward(ctx, bias, act, zero_row0, out_width, *ops):
        n = len(ops) // 2
        As, Ws = ops[:n], ops[n:]
        out = None
        M, N = As[0].shape[0], Ws[0].shape[0]
        epilogue = bias is not None or act != ACT_NONE or zero_row0
        # Fewer output tiles than SMs: a GEMM with the fused epilogue can only hand out whole-K tiles (24 CTAs at
        # 3 k rows: 62 us measured), a plain one is cut along K over all SMs (20 us); the epilogue then runs as one
        # element-wise pass (3 us).  Large products keep the fused epilogue.
        two_step = epilogue and N % 4 == 0 and ((M + 127) // 128) * ((N + 255) // 256) < 148
        for i, (A, W) in enumerate(zip(As, Ws)):
            last = (i == n - 1) and not two_step
            out = gemm_nt(A, W, out=out, beta=0.0 if i == 0 else 1.0, bias=bias if last else None,
                          act=act if last else ACT_NONE, zero_row0=zero_row0 and last, out_width=out_width)
        if two_step:
            call("ggpm_bias_act", fptr(out), _row_stride(out), M, N, None if bias is None else fptr(bias), act,
                 int(zero_row0), stream_ptr(out.device))
        ctx.act, ctx.zero_row0, ctx.n, ctx.N = act, zero_row0, n, Ws[0].shape[0]
        ctx.has_bias = bias is not None
        ctx.save_for_backward(out, *ops)
        return out

    @staticmethod
    def backward(ctx, dY):
        out, *ops = ctx.saved_tensors
        n, N = ctx.n, ctx.N
        As, Ws = ops[:n], ops[n:]
        dY = _kmajor(dY)
        M = out.shape[0]
        dZ = torch.empty(M, N, dtype=torch.float32, device=out.device)
        call("ggpm_act_bwd", fptr(dY), _row_stride(dY), fptr(out), _row_stride(out), fptr(dZ), N, M, N, ctx.act,
             int(ctx.zero_row0), stream_ptr(out.device))
        dbias = None
        if ctx.has_bias and ctx.needs_input_grad[0]:
            dbias = torch.empty(N, dtype=torch.float32, device=out.device)
            call("ggpm_colsum", fptr(dZ), N, M, N, fptr(dbias), 0, stream_ptr(out.device))
        grads: List[Optional[torch.Tensor]] = []
        for i, (A, W) in enumerate(zip(As, Ws)):
            grads.append(gemm_nn(dZ, W) if ctx.needs_input_grad[4 + i] else None)
        for i, (A, W) in enumerate(zip(As, Ws)):
            grads.append(gemm_tn(dZ, A) if ctx.needs_input_grad[4 + n + i] else None)
        return (dbias, None, None, None, *grads)


def linear_act(As: Sequence[torch.Tensor], Ws: Sequence[torch.Tensor], bias, act=ACT_NONE, zero_row0=False,
               out_width=None):
    return LinearAct.apply(bias, act, zero_row0, out_width, *As, *Ws)


class SplitPad(torch.autograd.Function):
    """W [N, K], bias [N] -> column blocks of W zero-padded to [Np, kp_i] and bias_p [Np], one launch each way
    (ggpm_split_pad / ggpm_split_unpad)."""

    @staticmethod
    def forward(ctx, W, bias, Np, cols, kps):
        import ctypes
        W = W.contiguous()
        N, K = W.shape
        dev = W.device
        outs = [torch.empty(Np, kp, dtype=torch.float32, device=dev) for kp in kps]
        bias_p = None if bias is None else torch.empty(Np, dtype=torch.float32, device=dev)
        n = len(kps)
        IA = ctypes.c_int * n
        ctx.meta = (N, K, Np, n, [c[0] for c in cols], [c[1] for c in cols], list(kps), bias is not None)
        call("ggpm_split_pad", fptr(W), K, N, K, Np, n, IA(*[c[0] for c in cols]), IA(*[c[1] for c in cols]), IA(*kps),
             (ctypes.c_void_p * n)(*[o.data_ptr() for o in outs]), None if bias is None else fptr(bias.contiguous()),
             None if bias_p is None else fptr(bias_p), stream_ptr(dev))
        if bias_p is None:
            bias_p = torch.zeros(0, device=dev)
            ctx.mark_non_differentiable(bias_p)
        return (bias_p, *outs)

    @staticmethod
    def backward(ctx, dbias_p, *douts):
        import ctypes
        N, K, Np, n, c0, c1, kps, has_bias = ctx.meta
        dev = next(d for d in (dbias_p,) + douts if d is not None).device
        dW = torch.empty(N, K, dtype=torch.float32, device=dev)
        dbias = torch.empty(N, dtype=torch.float32, device=dev) if has_bias else None
        douts = [None if d is None else d.contiguous() for d in douts]
        IA = ctypes.c_int * n
        call("ggpm_split_unpad", fptr(dW), N, K, n, IA(*c0), IA(*c1), IA(*kps),
             (ctypes.c_void_p * n)(*[None if d is None else d.data_ptr() for d in douts]),
             None if (dbias_p is None or not has_bias) else fptr(dbias_p.contiguous()), None if dbias is None else fptr(dbias),
             stream_ptr(dev))
        return dW, dbias, None, None, None


def linear_act_split(As: Sequence[torch.Tensor], W: torch.Tensor, widths: Sequence[int], bias, act=ACT_NONE,
                     zero_row0=False):
    """act([A_0[:, :k_0] | A_1[:, :k_1] | ...] @ W^T + b) without building the concatenation: W [N, sum k_i] is cut into
    its column blocks and zero-padded (rows to Np = N rounded up to 32, each block to a multiple of 4 columns) by ONE
    kernel; A_i is read in place at the padded width (its extra columns meet zero weights).  Returns [M, Np]."""
    import torch.nn.functional as F_
    N = W.shape[0]
    Np = round_up(N, 32)
    cols, kps, As_p, c = [], [], [], 0
    for A, k in zip(As, widths):
        kp = round_up(k, 4)
        cols.append((c, c + k))
        kps.append(kp)
        c += k
        A = A[:, :kp] if A.shape[1] >= kp else F_.pad(A[:, :k], (0, kp - k))
        if A.stride(1) != 1 or _row_stride(A) % 4 != 0 or A.data_ptr() % 16 != 0:
            A = A.contiguous()
        As_p.append(A)
    if c != W.shape[1]:
        raise ValueError("linear_act_split: widths sum to %d, weight has %d columns" % (c, W.shape[1]))
    bias_p, *Ws_p = SplitPad.apply(W, bias, Np, cols, kps)
    return LinearAct.apply(bias_p if bias is not None else None, act, zero_row0, None, *As_p, *Ws_p)


def linear_act_padded(As: Sequence[torch.Tensor], Ws: Sequence[torch.Tensor], bias, act=ACT_NONE, zero_row0=False):
    """linear_act with every operand brought to tensor-core friendly shapes: each weight piece W_i [N, K_i] is
    zero-padded to [Np, Kp_i] (Np = N rounded up to 32, Kp_i = K_i rounded up to 4) and A_i is used at width Kp_i
    (its extra columns meet zero weight columns, so only their finiteness matters).  Returns [M, Np]; columns
    >= N are act(0) = 0 for the activations on this path (ReLU, tanh, none).  The padding is autograd-tracked
    torch, so parameter gradients come back at the reference shapes."""
    import torch.nn.functional as F_
    N = Ws[0].shape[0]
    Np = round_up(N, 32)
    As_p, Ws_p = [], []
    for A, W in zip(As, Ws):
        K = W.shape[1]
        Kp = round_up(K, 4)
        if A.shape[1] >= Kp:
            A = A[:, :Kp]
        else:
            A = F_.pad(A[:, :K], (0, Kp - K))
        if A.stride(1) != 1 or _row_stride(A) % 4 != 0 or A.data_ptr() % 16 != 0:
            A = A.contiguous()
        As_p.append(A)
        Ws_p.append(F_.pad(W, (0, Kp - K, 0, Np - N)))
    bias_p = None if bias is None else F_.pad(bias, (0, Np - N))
    return LinearAct.apply(bias_p, act, zero_row0, None, *As_p, *Ws_p)


# ----------------------------------------------------------------------------------------------
# message passing
# ----------------------------------------------------------------------------------------------
class MessageGraph:
    """Index side of one message-passing layer, prepared once per batch:
    b (bgraph CSR), src/k1/k2 (int32 per message: source node, one-hot column ids), out-edge lists."""

    def __init__(self, bgraph_csr: CSR, src: torch.Tensor, k1: torch.Tensor, k2: Optional[torch.Tensor],
                 n_nodes: int, n_table_rows: int):
        self.b = bgraph_csr
        self.src, self.k1, self.k2 = src, k1, k2
        self.n_nodes, self.n_table_rows = n_nodes, n_table_rows
        self._by_src = None

    @property
    def E1(self):
        return self.b.rows

    @property
    def by_src(self) -> CSR:
        """for every node, the messages it emits (transpose of the [E1, 1] padded tensor `src`)"""
        if self._by_src is None:
            self._by_src = csr_from_padded(self.src.view(-1, 1), self.n_nodes).T
        return self._by_src


_BWD_PAIRS = None


def _bwd_pairs() -> bool:
    global _BWD_PAIRS
    if _BWD_PAIRS is None:
        import os
        _BWD_PAIRS = os.environ.get("GGPM_BWD_PAIRS", "0") == "1"
    return _BWD_PAIRS


def _empty(*shape, device):
    return torch.empty(*shape, dtype=torch.float32, device=device)


class LSTMMessages(torch.autograd.Function):
    """`depth` rounds of the graph LSTM (ggpm/rnn.py:85-108) over a message graph.
    phi_x [N1, F]: node features feeding the message inputs; Wx [4Hp, F], T [ntab, 4Hp], U4 = [Ug; Uf] [4Hp, Hp]:
    packed gate weights (ggpm_b200.packing).  Returns (h, c): [E1, Hp] final states."""

    @staticmethod
    def forward(ctx, phi_x, Wx, T, U4, mg: MessageGraph, depth: int, H: int):
        Hp = U4.shape[1]
        U4 = U4.contiguous()
        Ug, Uf = U4[:3 * Hp], U4[3 * Hp:]
        E1, dev = mg.E1, Uf.device
        keep = any(ctx.needs_input_grad)
        Yn = gemm_nt(phi_x, Wx)                                    # [N1, 4Hp]
        slots = depth if keep else min(depth, 2)
        # with the fp16-plane path the backward reads h only through hist_h16: fp32 h needs two rotating slots
        planes = keep and (_lib.backend().ggpm_lstm_plan(Hp, depth) & 2) != 0
        rotate = planes and depth > 2
        hist_h = _empty(2 if rotate else slots, E1, Hp, device=dev)
        hist_c = _empty(slots, E1, Hp, device=dev)
        hist_gp = _empty(slots, E1, 4 * Hp, device=dev)
        scratch = _empty(E1, Hp, device=dev)
        w_split = _empty(2, 4 * Hp, Hp, device=dev)               # weight planes (tf32 or fp16) + f16x3 engine scalars
        # fp16 hi / lo planes of every h^t (f16x3 engine): the TMA-fed operand of the weight-gradient GEMM
        hist_h16 = torch.empty(depth, 2, E1, Hp, dtype=torch.float16, device=dev) if planes else None
        T = T.contiguous()
        call("ggpm_lstm_fwd", E1, H, Hp, depth, iptr(mg.b.row_ptr), iptr(mg.b.col), fptr(Yn), iptr(mg.src), fptr(T),
             iptr(mg.k1), None if mg.k2 is None else iptr(mg.k2), fptr(Ug), fptr(Uf), fptr(hist_h), fptr(hist_c),
             fptr(hist_gp), fptr(scratch), fptr(w_split), None if hist_h16 is None else hist_h16.data_ptr(), int(keep),
             int(rotate), stream_ptr(dev))
        last = (depth - 1) if keep else ((depth - 1) & 1)
        h_last = ((depth - 1) & 1) if rotate else last
        ctx.rotate = rotate
        ctx.mg, ctx.depth, ctx.H, ctx.Hp = mg, depth, H, Hp
        ctx.hist = (hist_h, hist_c, hist_gp, Yn, hist_h16) if keep else None
        ctx.save_for_backward(phi_x, Wx, T, U4)
        # fresh tensors, not views of the history the backward pass reads: an in-place edit of a returned state by the
        # caller must not be able to corrupt the saved activations (autograd's version counters do not cover ctx.hist)
        return hist_h[h_last].clone(), hist_c[last].clone()

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dh, dc):
        phi_x, Wx, T, U4 = ctx.saved_tensors
        mg, depth, H, Hp = ctx.mg, ctx.depth, ctx.H, ctx.Hp
        Ug, Uf = U4[:3 * Hp], U4[3 * Hp:]
        if ctx.hist is None:
            raise RuntimeError("LSTMMessages: the saved activations were consumed by a previous backward pass "
                               "(the library back-propagates in place; run the forward again)")
        hist_h, hist_c, hist_gp, Yn, hist_h16 = ctx.hist
        ctx.hist = None
        E1, dev = mg.E1, Uf.device
        dH = (torch.zeros(E1, Hp, dtype=torch.float32, device=dev) if dh is None
              else dh.clone(memory_format=torch.contiguous_format))
        dCf = None if dc is None else dc.contiguous()       # gradient of the returned cell state (ggpm/rnn.py:108)
        z_planes = None if hist_h16 is None else torch.empty(2, E1, 4 * Hp, dtype=torch.float16, device=dev)
        dX = _empty(E1, 4 * Hp, device=dev)
        dU4 = torch.zeros(4 * Hp, Hp, dtype=torch.float32, device=dev)     # [dUg; dUf] contiguous (single-launch wgrad)
        dUg, dUf = dU4[:3 * Hp], dU4[3 * Hp:]
        dC, dP, dS = _empty(2, E1, Hp, device=dev), _empty(E1, Hp, device=dev), _empty(E1, Hp, device=dev)
        S, dHn = _empty(E1, Hp, device=dev), _empty(E1, Hp, device=dev)
        bt = mg.b.T
        # per-pair stores + reverse-CSR gathers instead of atomics (fp16-plane path; see include/ggpm_b200.h)
        bt_pos = getattr(bt, "pos", None)
        # Opt-in (GGPM_BWD_PAIRS=1): measured neutral to slightly slower than the vector reductions (cfg4 137.7 vs 136.2 ms,
        # cfg2 12.28 vs 12.18 ms: the stores + gathers cost what the zeroing launch did) — its value is a cell kernel without atomics
        # (the weight-gradient GEMM's split-K partial sums still meet in fp32 reduce-adds).
        pair_rows = mg.b.col.numel() if (_bwd_pairs() and z_planes is not None and bt_pos is not None and depth > 1) else 0
        pair_buf = _empty(3, pair_rows, Hp, device=dev) if pair_rows else None
        call("ggpm_lstm_bwd", E1, H, Hp, depth, iptr(mg.b.row_ptr), iptr(mg.b.col), iptr(bt.row_ptr), iptr(bt.col),
             fptr(Yn), iptr(mg.src), fptr(T), iptr(mg.k1), None if mg.k2 is None else iptr(mg.k2), fptr(Ug), fptr(Uf),
             fptr(hist_h), fptr(hist_c), fptr(hist_gp), fptr(dH), None if dCf is None else fptr(dCf), fptr(dX), fptr(dUg),
             fptr(dUf), None, fptr(dC),
             fptr(dP), fptr(dS), fptr(S), fptr(dHn), fptr(_empty(2, Hp, 4 * Hp, device=dev)),
             None if hist_h16 is None else hist_h16.data_ptr(), None if z_planes is None else z_planes.data_ptr(),
             int(ctx.rotate), None if pair_buf is None else iptr(bt_pos), None if pair_buf is None else fptr(pair_buf),
             pair_rows, stream_ptr(dev))
        dphi, dWx, dT = _xproj_backward(ctx, dX, phi_x, Wx, T, mg)
        return dphi, dWx, dT, dU4, None, None, None


def _xproj_backward(ctx, dX, phi_x, Wx, T, mg: MessageGraph):
    """dX [E1, G*Hp] -> gradients of the hoisted input projection: node features, packed x-weights, tables"""
    dev = dX.device
    dphi = dWx = dT = None
    if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
        dYn = _gather_sum_raw(dX, mg.by_src, dX.shape[1])                 # [N1, G*Hp]
        if ctx.needs_input_grad[0]:
            dphi = gemm_nn(dYn, Wx)
        if ctx.needs_input_grad[1]:
            dWx = gemm_tn(dYn, phi_x)
    if ctx.needs_input_grad[2]:
        dT = torch.zeros_like(T)
        call("ggpm_table_wgrad", fptr(dX), dX.shape[1], dX.shape[0], dX.shape[1], iptr(mg.k1),
             None if mg.k2 is None else iptr(mg.k2), fptr(dT), dT.shape[1], dT.shape[0], stream_ptr(dev))
    return dphi, dWx, dT


class GRUMessages(torch.autograd.Function):
    """`depth` rounds of the graph GRU (ggpm/rnn.py:25-50).  Wx [3Hp, F], T [ntab, 3Hp] packed [z|h|r];
    Uz, Ur, Uh [Hp, Hp]; br [Hp].  Returns h [E1, Hp]."""

    @staticmethod
    def forward(ctx, phi_x, Wx, T, Uz, Ur, br, Uh, mg: MessageGraph, depth: int, H: int):
        Hp = Uz.shape[0]
        E1, dev = mg.E1, Uz.device
        keep = any(ctx.needs_input_grad)
        Yn = gemm_nt(phi_x, Wx)                                    # [N1, 3Hp]
        hist_h = _empty(depth + 1 if keep else 2, E1, Hp, device=dev)
        hist_aux = _empty(depth if keep else 1, E1, 5 * Hp, device=dev)
        T, Uz, Ur, br, Uh = (t.contiguous() for t in (T, Uz, Ur, br, Uh))
        call("ggpm_gru_fwd", E1, H, Hp, depth, iptr(mg.b.row_ptr), iptr(mg.b.col), fptr(Yn), iptr(mg.src), fptr(T),
             iptr(mg.k1), None if mg.k2 is None else iptr(mg.k2), fptr(Uz), fptr(Ur), fptr(br), fptr(Uh), fptr(hist_h),
             fptr(hist_aux), int(keep), stream_ptr(dev))
        ctx.mg, ctx.depth, ctx.H, ctx.Hp = mg, depth, H, Hp
        ctx.hist = (hist_h, hist_aux, Yn) if keep else None
        ctx.save_for_backward(phi_x, Wx, T, Uz, Ur, br, Uh)
        return hist_h[depth if keep else (depth & 1)]

    @staticmethod
    def backward(ctx, dh):
        phi_x, Wx, T, Uz, Ur, br, Uh = ctx.saved_tensors
        mg, depth, H, Hp = ctx.mg, ctx.depth, ctx.H, ctx.Hp
        hist_h, hist_aux, Yn = ctx.hist
        E1, dev = mg.E1, Uz.device
        dH = dh.clone(memory_format=torch.contiguous_format)
        dX = _empty(E1, 3 * Hp, device=dev)
        dUz, dUr, dUh, dbr = torch.zeros_like(Uz), torch.zeros_like(Ur), torch.zeros_like(Uh), torch.zeros_like(br)
        dA = _empty(E1, 2 * Hp, device=dev)
        dS, dSg, dPr, dHn = (_empty(E1, Hp, device=dev) for _ in range(4))
        call("ggpm_gru_bwd", E1, H, Hp, depth, iptr(mg.b.row_ptr), iptr(mg.b.col), fptr(Yn), iptr(mg.src), fptr(T),
             iptr(mg.k1), None if mg.k2 is None else iptr(mg.k2), fptr(Uz), fptr(Ur), fptr(Uh), fptr(hist_h),
             fptr(hist_aux), fptr(dH), fptr(dX), fptr(dUz), fptr(dUr), fptr(dbr), fptr(dUh), fptr(dA), fptr(dS),
             fptr(dSg), fptr(dPr), fptr(dHn), stream_ptr(dev))
        dphi, dWx, dT = _xproj_backward(ctx, dX, phi_x, Wx, T, mg)
        return dphi, dWx, dT, dUz, dUr, dbr, dUh, None, None, None


# ----------------------------------------------------------------------------------------------
# incremental ("sparse") message passing of the decoder's encoder — SURVEY 8(f) rank 1
# ----------------------------------------------------------------------------------------------
def _rows(t: torch.Tensor) -> torch.Tensor:
    """fp32 2-D with unit column stride (row stride free)"""
    if t.dim() != 2 or t.dtype != torch.float32:
        raise ValueError("expected a 2-D float32 tensor, got %s %s" % (tuple(t.shape), t.dtype))
    return t if (t.stride(1) == 1 or t.shape[1] == 1) else t.contiguous()


def row_ids(index: torch.Tensor, n_rows: int, check: bool = False) -> torch.Tensor:
    """int64 / int32 row ids -> contiguous int32 (any shape, flattened by the callers)"""
    _lib.require_device(index, "index tensor")
    if index.dtype not in (torch.int64, torch.int32):
        raise ValueError("index tensors must be int64 or int32, got %s" % index.dtype)
    if check and index.numel() and (int(index.min()) < 0 or int(index.max()) >= n_rows):
        raise IndexError("row index out of range [0, %d)" % n_rows)
    return index.to(torch.int32).contiguous()


def rows_gather(table, idx, M, K, pairs=True, total=True, pad_to=None):
    """pairs [M*K, Wp] = table[idx], total [M, Wp] = sum over each row's K slots (one kernel); Wp = pad_to (default:
    the table width), columns beyond the table width are zero"""
    table = _rows(table)
    W = table.shape[1]
    Wp = W if pad_to is None else pad_to
    dev = table.device
    P_ = torch.empty(M * K, Wp, dtype=torch.float32, device=dev) if pairs else None
    S_ = torch.empty(M, Wp, dtype=torch.float32, device=dev) if total else None
    call("ggpm_rows_gather", fptr(table), _row_stride(table), iptr(idx), M, K, W, Wp, None if P_ is None else fptr(P_), Wp,
         None if S_ is None else fptr(S_), Wp, stream_ptr(dev))
    return P_, S_


def rows_scatter_(table, idx, src=None):
    """table[idx[m]] = src[m, :table width] (zero rows when src is None), in place"""
    M = idx.numel()
    if M == 0:
        return table
    if src is not None:
        src = _rows(src)
    call("ggpm_rows_scatter", fptr(table), _row_stride(table), iptr(idx), M, table.shape[1],
         None if src is None else fptr(src), 0 if src is None else _row_stride(src), stream_ptr(table.device))
    return table


def rows_scatter_add_(table, idx, M, K, src=None, bcast=None):
    """table[idx[p]] += src[p] + bcast[p // K], in place"""
    call("ggpm_rows_scatter_add", fptr(table), _row_stride(table), iptr(idx), M, K, table.shape[1],
         None if src is None else fptr(src), 0 if src is None else _row_stride(src),
         None if bcast is None else fptr(bcast), 0 if bcast is None else _row_stride(bcast), stream_ptr(table.device))
    return table


class IndexScatter(torch.autograd.Function):
    """index_scatter(sub_data, all_data, index) of ggpm/nnutils.py:124-128: a copy of all_data whose rows `index` are
    replaced by sub_data — one table copy + one row scatter instead of six full-table passes."""

    @staticmethod
    def forward(ctx, sub, table, idx):
        out = table.contiguous().clone()
        rows_scatter_(out, idx, sub)
        ctx.idx = idx
        return out

    @staticmethod
    def backward(ctx, dout):
        idx = ctx.idx
        dout = _rows(dout)
        d_sub = d_tab = None
        if ctx.needs_input_grad[0]:
            d_sub, _ = rows_gather(dout, idx, idx.numel(), 1, pairs=True, total=False)
        if ctx.needs_input_grad[1]:
            d_tab = rows_scatter_(dout.contiguous().clone(), idx, None)
        return d_sub, d_tab, None


class ScatterIntoZeros(torch.autograd.Function):
    """index_scatter(sub, zeros(R, W), index) — the node_buf pattern of ggpm/encoder.py:176-178, 224-226 — without the
    table copies: one memset + one row scatter forward, one row gather backward."""

    @staticmethod
    def forward(ctx, sub, idx, R):
        out = torch.zeros(R, sub.shape[1], dtype=torch.float32, device=sub.device)
        rows_scatter_(out, idx, sub)
        ctx.idx = idx
        return out

    @staticmethod
    def backward(ctx, dout):
        return rows_gather(_rows(dout), ctx.idx, ctx.idx.numel(), 1, total=False)[0], None, None


def scatter_into_zeros(sub_data, index, n_rows):
    if sub_data.shape[0] != index.numel():
        raise ValueError("scatter_into_zeros: %d rows for %d indices" % (sub_data.shape[0], index.numel()))
    if index.numel() == 0:
        return sub_data.new_zeros(n_rows, sub_data.shape[1])
    return ScatterIntoZeros.apply(_rows(sub_data), row_ids(index.reshape(-1), n_rows), n_rows)


def index_scatter(sub_data, all_data, index):
    if sub_data.shape[0] != index.numel() or sub_data.shape[1] != all_data.shape[1]:
        raise ValueError("index_scatter: sub_data %s does not match index %d / table %s"
                         % (tuple(sub_data.shape), index.numel(), tuple(all_data.shape)))
    if index.numel() == 0:
        return all_data * 1.0
    return IndexScatter.apply(_rows(sub_data), all_data, row_ids(index.reshape(-1), all_data.shape[0]))


def _stack(depth, rows, width, dev):
    return torch.empty(depth, rows, width, dtype=torch.float32, device=dev)


class SparseLSTM(torch.autograd.Function):
    """LSTM.sparse_forward (ggpm/rnn.py:110-121) as one autograd node and ONE library call per direction
    (ggpm_sparse_lstm_fwd / _bwd sequence all the rounds): rows `sub` of the tables (h, c) [R, H] are zeroed and
    recomputed for `depth` rounds from the K padded neighbour slots `nbr` [M*K] read out of the whole table; all other
    rows pass through.  x [M, Fp] are the (embedded) inputs of the sub-messages; Wx [4Hp, Fp], T [1, 4Hp] (biases) and
    U4 [4Hp, Hp] = (U_i; U_o; U_u; U_f) come from packing.pack_lstm (gate order i, o, u, f; zero padded).  Between
    gather and scatter everything lives at the padded width Hp, so the H x H products run on the tensor-core engine.
    Per call: one table copy each way; per round: M*K gathered rows, two small GEMMs, one cell kernel, M scattered
    rows."""

    @staticmethod
    def forward(ctx, h, c, x, Wx, T, U4, sub, nbr, K, depth):
        M, Hp, H = sub.numel(), U4.shape[1], h.shape[1]
        dev = h.device
        Xg = gemm_nt(x, Wx, bias=T)
        hw, cw = h.contiguous().clone(), c.contiguous().clone()
        S, gates = _stack(depth, M, Hp, dev), _stack(depth, M, 4 * Hp, dev)
        Hn, Cn, fs = (_stack(depth, M * K, Hp, dev) for _ in range(3))
        A = torch.empty(M, 6 * Hp, dtype=torch.float32, device=dev)              # A [M, 4Hp] | h_sub | c_sub
        call("ggpm_sparse_lstm_fwd", M, K, H, Hp, depth, fptr(hw), fptr(cw), H, iptr(sub), iptr(nbr), fptr(Xg), fptr(U4),
             fptr(S), fptr(Hn), fptr(Cn), fptr(gates), fptr(fs), A.data_ptr(), A.data_ptr() + 16 * M * Hp,
             A.data_ptr() + 20 * M * Hp, stream_ptr(dev))
        ctx.dims = (M, K, H, Hp, depth)
        ctx.sub, ctx.nbr = sub, nbr
        ctx.save_for_backward(x, Wx, U4, S, Hn, Cn, gates, fs)
        return hw, cw

    @staticmethod
    def backward(ctx, dh, dc):
        M, K, H, Hp, depth = ctx.dims
        sub, nbr = ctx.sub, ctx.nbr
        x, Wx, U4, S, Hn, Cn, gates, fs = ctx.saved_tensors
        dev = x.device
        dHw = dh.contiguous().clone()
        dCw = dc.contiguous().clone()
        dXg = torch.empty(M, 4 * Hp, dtype=torch.float32, device=dev)
        dU4 = torch.empty_like(U4)
        rowbuf = torch.empty(M, 7 * Hp, dtype=torch.float32, device=dev)         # dA [M, 4Hp] | dh_sub | dc_sub | dS
        pairbuf = torch.empty(3, M * K, Hp, dtype=torch.float32, device=dev)     # dPf | dCn | dHn
        r, p, f4 = rowbuf.data_ptr(), pairbuf.data_ptr(), 4 * M * Hp
        call("ggpm_sparse_lstm_bwd", M, K, H, Hp, depth, fptr(dHw), fptr(dCw), H, iptr(sub), iptr(nbr), fptr(U4), fptr(S),
             fptr(Hn), fptr(Cn), fptr(gates), fptr(fs), fptr(dXg), fptr(dU4), r, p, p + f4 * K, r + 4 * f4, r + 5 * f4,
             r + 6 * f4, p + 2 * f4 * K, stream_ptr(dev))
        dx = gemm_nn(dXg, Wx) if ctx.needs_input_grad[2] else None
        dWx = gemm_tn(dXg, x)
        dT = torch.empty(1, 4 * Hp, dtype=torch.float32, device=dev)
        call("ggpm_colsum", fptr(dXg), 4 * Hp, M, 4 * Hp, fptr(dT), 0, stream_ptr(dev))
        return dHw, dCw, dx, dWx, dT, dU4, None, None, None, None


class SparseGRU(torch.autograd.Function):
    """GRU.sparse_forward (ggpm/rnn.py:52-59), one library call per direction (ggpm_sparse_gru_fwd / _bwd).  Operands
    from packing.pack_gru: Wx [3Hp, Fp], T [1, 3Hp] in order (z, h~, r) — T carries the W_z / W_h biases, br [Hp] is
    U_r's (W_r has none); Uz, Ur, Uh [Hp, Hp]."""

    @staticmethod
    def forward(ctx, h, x, Wx, T, Uz, Ur, br, Uh, sub, nbr, K, depth):
        M, Hp, H = sub.numel(), Uz.shape[1], h.shape[1]
        dev = h.device
        Xg = gemm_nt(x, Wx, bias=torch.cat([T[0, :2 * Hp], br]))    # the r pre-activation carries U_r's bias
        hw = h.contiguous().clone()
        Uz, Ur, Uh = Uz.contiguous(), Ur.contiguous(), Uh.contiguous()
        S, Sg, zh = _stack(depth, M, Hp, dev), _stack(depth, M, Hp, dev), _stack(depth, M, 2 * Hp, dev)
        Hn, rs = _stack(depth, M * K, Hp, dev), _stack(depth, M * K, Hp, dev)
        A = torch.empty(M, 3 * Hp, dtype=torch.float32, device=dev)              # A [M, 2Hp] | h_sub
        call("ggpm_sparse_gru_fwd", M, K, H, Hp, depth, fptr(hw), H, iptr(sub), iptr(nbr), fptr(Xg), fptr(Uz), fptr(Ur),
             fptr(Uh), fptr(S), fptr(Hn), fptr(rs), fptr(Sg), fptr(zh), A.data_ptr(), A.data_ptr() + 8 * M * Hp,
             stream_ptr(dev))
        ctx.dims = (M, K, H, Hp, depth)
        ctx.sub, ctx.nbr = sub, nbr
        ctx.save_for_backward(x, Wx, Uz, Ur, Uh, S, Hn, rs, Sg, zh)
        return hw

    @staticmethod
    def backward(ctx, dh):
        M, K, H, Hp, depth = ctx.dims
        sub, nbr = ctx.sub, ctx.nbr
        x, Wx, Uz, Ur, Uh, S, Hn, rs, Sg, zh = ctx.saved_tensors
        dev = x.device
        dHw = dh.contiguous().clone()
        dXg = torch.empty(M, 3 * Hp, dtype=torch.float32, device=dev)
        dUz, dUr, dUh = torch.empty_like(Uz), torch.empty_like(Ur), torch.empty_like(Uh)
        rowbuf = torch.empty(M, 5 * Hp, dtype=torch.float32, device=dev)         # dA [M, 2Hp] | dS | dSg | dh_sub
        pairbuf = torch.empty(2, M * K, Hp, dtype=torch.float32, device=dev)     # dPr | dHn
        r, p, f4 = rowbuf.data_ptr(), pairbuf.data_ptr(), 4 * M * Hp
        call("ggpm_sparse_gru_bwd", M, K, H, Hp, depth, fptr(dHw), H, iptr(sub), iptr(nbr), fptr(Uz), fptr(Ur), fptr(Uh),
             fptr(S), fptr(Hn), fptr(rs), fptr(Sg), fptr(zh), fptr(dXg), fptr(dUz), fptr(dUr), fptr(dUh), r, r + 2 * f4,
             r + 3 * f4, p, p + f4 * K, r + 4 * f4, stream_ptr(dev))
        dx = gemm_nn(dXg, Wx) if ctx.needs_input_grad[1] else None
        dWx = gemm_tn(dXg, x)
        dT = torch.empty(1, 3 * Hp, dtype=torch.float32, device=dev)
        call("ggpm_colsum", fptr(dXg), 3 * Hp, M, 3 * Hp, fptr(dT), 0, stream_ptr(dev))
        dbr = dT[0, 2 * Hp:].clone()
        return dHw, dx, dWx, dT, dUz, dUr, dbr, dUh, None, None, None, None


class SlotGatherSum(torch.autograd.Function):
    """out[m] = sum_k table[idx[m*K + k]] over ALL K slots (slot value 0 reads row 0): index_select_ND(table, 0,
    index).sum(1) (ggpm/encoder.py:171-172, 217) and, with K = 1, table.index_select(0, index) (:220, 229)."""

    @staticmethod
    def forward(ctx, table, idx, M, K):
        ctx.idx, ctx.dims = idx, (M, K, table.shape[0])
        return rows_gather(table, idx, M, K, pairs=False)[1]

    @staticmethod
    def backward(ctx, dout):
        M, K, R = ctx.dims
        dt = torch.zeros(R, dout.shape[1], dtype=torch.float32, device=dout.device)
        rows_scatter_add_(dt, ctx.idx, M, K, bcast=_rows(dout))
        return dt, None, None, None


def slot_gather_sum(table: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    """index [M, K] (or [M]) of row ids into table [R, W] -> [M, W]"""
    M = index.shape[0]
    K = 1 if index.dim() == 1 else index.shape[1]
    if M == 0 or K == 0:
        return table.new_zeros(M, table.shape[1])
    return SlotGatherSum.apply(_rows(table), row_ids(index.reshape(-1), table.shape[0]), M, K)
