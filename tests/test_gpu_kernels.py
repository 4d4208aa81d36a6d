"""B200: every CUDA kernel, called through the C ABI, against its torch twin (tests/_proto.py, fp64) or the
numpy index oracle (oracle/tensorize_oracle.py) on identical seeded inputs.  Integer/index work is compared
bit-exactly; fp32 arithmetic within 1e-5 relative of the fp64 twin."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _proto as P                                                     # noqa: E402
from conftest import load_golden                                       # noqa: E402
from ggpm_b200 import _lib, functional as Fn, synth                    # noqa: E402
from oracle import tensorize_oracle as to                              # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30)


@pytest.fixture(params=[1, 2, 0], ids=["simt", "tf32x3", "auto"])
def engine(request):
    _lib.set_gemm_engine(request.param)
    yield request.param
    _lib.set_gemm_engine(0)


def random_padded(rng, rows, width, n_src, density=0.5, mid_zeros=True):
    idx = rng.integers(1, n_src, size=(rows, width))
    keep = rng.random((rows, width)) < density
    if not mid_zeros:
        keep = np.sort(keep, axis=1)[:, ::-1]
    idx = idx * keep
    idx[0] = 0
    return idx.astype(np.int64)


# ---------------------------------------------------------------------------------------------------
# index construction: bit-exact
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rows,width,n_src,dtype", [(1, 1, 1, torch.int64), (2, 3, 2, torch.int64),
                                                     (1000, 6, 777, torch.int64), (5000, 9, 5000, torch.int32),
                                                     (300000, 5, 300000, torch.int64)])
def test_csr_from_padded_and_transpose_bit_exact(rows, width, n_src, dtype):
    rng = np.random.default_rng(rows + width)
    idx = random_padded(rng, rows, width, max(n_src, 2)) % n_src
    csr = Fn.csr_from_padded(torch.from_numpy(idx).to(dtype).to(DEV), n_src, check=True)
    rp, col = to.csr_from_padded_oracle(idx)
    assert np.array_equal(csr.row_ptr.cpu().numpy(), rp)
    nnz = int(rp[-1])
    assert np.array_equal(csr.col[:nnz].cpu().numpy(), col)
    tp, trow, tpos = to.csr_transpose_oracle(rp, col, n_src)
    t = csr.T
    assert np.array_equal(t.row_ptr.cpu().numpy(), tp)
    assert np.array_equal(t.col[:nnz].cpu().numpy(), trow)
    assert np.array_equal(t.pos[:nnz].cpu().numpy(), tpos)


def test_csr_of_golden_batch_and_bad_index():
    meta, z, tree, graph = load_golden("hier_lstm_cfg1")
    for idx, n_src in ((tree[2], tree[1].shape[0]), (tree[3], tree[1].shape[0]), (tree[4], graph[0].shape[0]),
                       (graph[2], graph[1].shape[0]), (graph[3], graph[1].shape[0])):
        csr = Fn.csr_from_padded(idx.to(DEV), n_src, check=True)
        rp, col = to.csr_from_padded_oracle(idx.numpy())
        assert np.array_equal(csr.row_ptr.cpu().numpy(), rp)
        assert np.array_equal(csr.col[:rp[-1]].cpu().numpy(), col)
    with pytest.raises(IndexError):
        Fn.csr_from_padded(torch.tensor([[0, 0], [5, 0]], device=DEV), 3, check=True)


# ---------------------------------------------------------------------------------------------------
# gather-sum
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rows,n_src,width,K", [(1, 1, 4, 1), (50, 40, 40, 3), (2000, 1500, 256, 6),
                                                 (3000, 3000, 512, 4), (700, 900, 2816, 5), (64, 64, 1024, 40)])
def test_gather_sum_forward_backward(rows, n_src, width, K):
    rng = np.random.default_rng(rows)
    idx = torch.from_numpy(random_padded(rng, rows, K, max(n_src, 2)) % n_src)
    table = torch.randn(n_src, width, dtype=torch.float64)
    table[0] = 0
    rp, col = P.csr_from_padded(idx)
    expect = P.gather_sum(table, rp, col)
    t = table.float().to(DEV).requires_grad_(True)
    csr = Fn.csr_from_padded(idx.to(DEV), n_src)
    out = Fn.gather_sum(t, csr)
    assert rel(out, expect) <= 2e-6
    cot = torch.randn(rows, width, dtype=torch.float64)
    out.backward(cot.float().to(DEV))
    assert rel(t.grad, P.scatter_rows(cot, rp, col, n_src)) <= 2e-6


def test_gather_sum_unaligned_table_is_repacked():
    idx = torch.tensor([[0, 0], [1, 2], [2, 0]])
    table = torch.randn(3, 10, device=DEV)[:, 1:8]           # 7 columns, misaligned base
    out = Fn.gather_sum(table, Fn.csr_from_padded(idx.to(DEV), 3))
    ref = torch.stack([torch.zeros(7, device=DEV), table[1] + table[2], table[2]])
    assert torch.allclose(out, ref, atol=1e-6)


# ---------------------------------------------------------------------------------------------------
# GEMM (all operand layouts, ragged sizes, epilogues)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K", [(1, 1, 1), (300, 250, 38), (129, 257, 250), (1000, 768, 256), (4097, 256, 256),
                                    (513, 1024, 512), (200, 704, 720), (130, 2112, 704)])
def test_gemm_nt_with_epilogue(engine, M, N, K):
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g, dtype=torch.float64)
    W = torch.randn(N, K, generator=g, dtype=torch.float64) / K ** 0.5
    b = torch.randn(N, generator=g, dtype=torch.float64)
    C0 = torch.randn(M, N, generator=g, dtype=torch.float64)
    for act, fn in ((0, lambda x: x), (1, torch.relu), (2, torch.tanh)):
        out = C0.float().to(DEV)
        Fn.gemm_nt(A.float().to(DEV), W.float().to(DEV), out=out, beta=1.0, bias=b.float().to(DEV), act=act,
                   zero_row0=True)
        ref = fn(A.float().double() @ W.float().double().t() + C0.float().double() + b.float().double())
        ref[0] = 0
        assert rel(out, ref) <= 1e-5, (act, rel(out, ref))


@pytest.mark.parametrize("M,N,K", [(4096, 256, 1000), (5000, 512, 520), (4200, 250, 538), (6000, 128, 388)])
def test_gemm_nt_long_k_runs_as_fp32_accumulated_slices(engine, M, N, K):
    """generic GEMMs with >= 4096 rows and K > 384 run as 256-deep K-slices (gemm_tc.cu launch_gemm_nt): every epilogue
    (accumulate, bias, activation, row-0 mask) gives the unsliced result, and the post-ReLU (non-negative) operand that
    makes the tensor core's round-toward-zero accumulation coherent stays at fp32-level error"""
    g = torch.Generator().manual_seed(M + K)
    A = torch.relu(torch.randn(M, K, generator=g, dtype=torch.float64))
    W = torch.randn(N, K, generator=g, dtype=torch.float64) / K ** 0.5
    b = torch.randn(N, generator=g, dtype=torch.float64) * 0.1
    C0 = torch.randn(M, N, generator=g, dtype=torch.float64)
    Af, Wf, bf = A.float().to(DEV), W.float().to(DEV), b.float().to(DEV)
    lin = Af.double().cpu() @ Wf.double().cpu().t()
    for beta, bias, act, z0 in ((0.0, None, 0, False), (1.0, bf, 1, True), (0.0, bf, 2, False), (1.0, None, 0, False)):
        out = C0.float().to(DEV)
        Fn.gemm_nt(Af, Wf, out=out, beta=beta, bias=bias, act=act, zero_row0=z0)
        ref = lin + beta * C0.float().double() + (0 if bias is None else bf.double().cpu())
        ref = (lambda x: x, torch.relu, torch.tanh)[act](ref)
        if z0:
            ref[0] = 0
        assert rel(out, ref) <= 1e-5, (beta, act, rel(out, ref))
    out = Fn.gemm_nt(Af, Wf)
    err = out.double().cpu() - lin
    assert float(err.pow(2).mean().sqrt() / lin.pow(2).mean().sqrt()) <= (1.2e-6 if engine != 1 else 6e-7)


def test_gemm_strided_views_and_padded_output(engine):
    g = torch.Generator().manual_seed(5)
    big = torch.randn(500, 300, generator=g).to(DEV)
    Wfull = torch.randn(250, 538, generator=g).to(DEV) * 0.05
    A = big[:, :250]                       # row stride 300
    W = Wfull[:, 288:]                     # row stride 538, misaligned base
    out = Fn.gemm_nt(A, W, out_width=256)
    assert out.shape == (500, 256) and float(out[:, 250:].abs().max()) == 0.0
    assert rel(out[:, :250], A.double() @ W.double().t()) <= 1e-5


@pytest.mark.parametrize("M,N,K", [(300, 250, 768), (1000, 256, 256), (77, 40, 1024)])
def test_gemm_nn_input_gradient(M, N, K):
    A, B = torch.randn(M, K).to(DEV), (torch.randn(K, N) / K ** 0.5).to(DEV)
    base = torch.randn(M, N).to(DEV)
    out = base.clone()
    Fn.gemm_nn(A, B, out=out, beta=1.0)
    assert rel(out, A.double() @ B.double() + base.double()) <= 1e-5
    assert rel(Fn.gemm_nn(A, B), A.double() @ B.double()) <= 1e-5


@pytest.mark.parametrize("K,M,N", [(1, 8, 8), (5000, 768, 256), (16001, 256, 250), (300, 38, 1024), (70000, 256, 256)])
def test_gemm_tn_weight_gradient_split_k(K, M, N):
    A, B = torch.randn(K, M).to(DEV), torch.randn(K, N).to(DEV)
    ref = A.double().t() @ B.double()
    assert rel(Fn.gemm_tn(A, B), ref) <= 1e-5
    acc = torch.ones(M, N, device=DEV)
    Fn.gemm_tn(A, B, out=acc, beta=1.0)
    assert rel(acc, ref + 1.0) <= 1e-5
    over = torch.full((M, N), 7.0, device=DEV)
    Fn.gemm_tn(A, B, out=over, beta=0.0)
    assert rel(over, ref) <= 1e-5


# ---------------------------------------------------------------------------------------------------
# tcgen05 GEMM with the gather-sum fused into the operand prologue, and the single-launch layer forms
# ---------------------------------------------------------------------------------------------------
def _rand_csr(rng, rows, n_src, width=4, density=0.5):
    idx = torch.from_numpy(random_padded(rng, rows, width, n_src, density)).to(DEV)
    return idx, Fn.csr_from_padded(idx, n_src)


def _gsum(table, idx):
    """index_select_ND(table, 0, idx).sum(1) in fp64 with zero entries skipped (row 0 of real tables is zero)"""
    t = table.double()
    g = t[idx.reshape(-1)].reshape(idx.shape + (t.shape[1],))
    return (g * (idx != 0).unsqueeze(-1)).sum(1)


@pytest.mark.parametrize("E,Hp,tile_n", [(300, 128, 0), (1000, 256, 0), (1000, 256, 128), (2500, 512, 0)])
def test_fused_gate_projection_forward(E, Hp, tile_n):
    """[G | P] = [gather_sum(h) Ug^T | h Uf^T] in one launch (ggpm/rnn.py:86-91)"""
    rng = np.random.default_rng(E + Hp)
    h = torch.randn(E, Hp, device=DEV)
    U4 = torch.randn(4 * Hp, Hp, device=DEV) / Hp ** 0.5
    idx, csr = _rand_csr(rng, E, E)
    out = torch.empty(E, 4 * Hp, device=DEV)
    Fn.gemm_fused(h, Hp, 1, U4, 1, Hp, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp, tile_n=tile_n)
    S = _gsum(h, idx)
    ref = torch.cat([S @ U4[:3 * Hp].double().t(), h.double() @ U4[3 * Hp:].double().t()], 1)
    assert rel(out, ref) <= 1e-5, rel(out, ref)


@pytest.mark.parametrize("E,Hp", [(300, 128), (1000, 256), (5000, 256)])
def test_fused_state_gradient(E, Hp):
    """dH = [gather^T(dA[:, :3Hp]) | dP] [Ug; Uf] in one launch (backward of ggpm/rnn.py:86-91)"""
    rng = np.random.default_rng(E)
    dA = torch.randn(E, 4 * Hp, device=DEV)
    dP = torch.randn(E, Hp, device=DEV)
    U4 = torch.randn(4 * Hp, Hp, device=DEV) / Hp ** 0.5
    idx, csr = _rand_csr(rng, E, E)
    bt = csr.T
    out = torch.empty(E, Hp, device=DEV)
    Fn.gemm_fused(dA, 4 * Hp, 1, U4, Hp, 1, out, E, Hp, 4 * Hp, a_csr=bt, split_dim=2, split_at=3 * Hp, A2=dP, lda2=Hp)
    dS = dA[:, :3 * Hp].double() @ U4[:3 * Hp].double()
    scat = torch.zeros(E, Hp, dtype=torch.float64, device=DEV)
    flat = idx.reshape(-1)
    rows = torch.arange(E, device=DEV).repeat_interleave(idx.shape[1])
    keep = flat != 0
    scat.index_add_(0, flat[keep], dS[rows[keep]])
    ref = scat + dP.double() @ U4[3 * Hp:].double()
    assert rel(out, ref) <= 1e-5, rel(out, ref)


@pytest.mark.parametrize("E,Hp", [(300, 128), (5000, 256), (40000, 256)])
def test_fused_weight_gradient(E, Hp):
    """[dUg; dUf] += [dA[:, :3Hp]^T gather_sum(h) ; dP^T h] in one split-K launch"""
    rng = np.random.default_rng(E + 1)
    dA = torch.randn(E, 4 * Hp, device=DEV)
    dP = torch.randn(E, Hp, device=DEV)
    h = torch.randn(E, Hp, device=DEV)
    idx, csr = _rand_csr(rng, E, E)
    out = torch.ones(4 * Hp, Hp, device=DEV)
    Fn.gemm_fused(dA, 1, 4 * Hp, h, Hp, 1, out, 4 * Hp, Hp, E, beta=1.0, b_csr=csr, split_dim=3, split_at=3 * Hp, A2=dP,
                  lda2=Hp, split_k=max(1, min(18, E // 256)))
    S = _gsum(h, idx)
    ref = torch.cat([dA[:, :3 * Hp].double().t() @ S, dP.double().t() @ h.double()], 0) + 1.0
    assert rel(out, ref) <= 1e-5, rel(out, ref)


# ---------------------------------------------------------------------------------------------------
# f16x3 engine (fp16 two-term split, tcgen05.mma.kind::f16)
# ---------------------------------------------------------------------------------------------------
def _max_deg(csr):
    rp = csr.row_ptr
    return (rp[1:] - rp[:-1]).max().to(torch.int32).reshape(1).contiguous()


@pytest.mark.parametrize("M,N,K,scale_a,scale_w", [(300, 128, 64, 1.0, 1.0), (1000, 1024, 256, 1.0, 0.05),
                                                   (4100, 512, 512, 1e-6, 30.0), (129, 256, 96, 1e4, 1e-3),
                                                   (20000, 1024, 256, 1.0, 0.06)])
def test_f16x3_plain_gemm(M, N, K, scale_a, scale_w):
    torch.manual_seed(M + N)
    A = torch.randn(M, K, device=DEV) * scale_a
    A[::7] *= 1e-4                                      # rows far below the tensor max keep absolute accuracy
    W = torch.randn(N, K, device=DEV) * scale_w
    planes, scal = Fn.f16_prepare_weights(W)
    amax = A.abs().max().reshape(1)
    out = torch.full((M, N), 7.0, device=DEV)
    Fn.gemm_f16x3(A, K, planes, scal, out, M, N, K, a_amax=amax)
    ref = A.double() @ W.double().t()
    assert rel(out, ref) <= 2e-6, rel(out, ref)
    assert float(scal[0]) == float(W.abs().max())


def test_f16x3_epilogue_and_transposed_planes():
    torch.manual_seed(5)
    M, N, K = 777, 384, 160
    A = torch.rand(M, K, device=DEV) * 2 - 1
    W = torch.randn(K, N, device=DEV) * 0.1              # B(k, n) = W[k, n]: planes of the transpose
    planes, scal = Fn.f16_prepare_weights(W, transpose=True)
    assert planes.shape == (2, N, K)
    bias = torch.randn(N, device=DEV)
    out = torch.randn(M, N, device=DEV)
    ref = torch.tanh(A.double() @ W.double() + out.double() + bias.double())
    ref[0] = 0
    Fn.gemm_f16x3(A, K, planes, scal, out, M, N, K, beta=1.0, bias=bias, act=_lib.ACT_TANH, zero_row0=True,
                  a_amax_const=1.0)
    assert rel(out, ref) <= 2e-6, rel(out, ref)


@pytest.mark.parametrize("E,Hp,tile_n", [(300, 128, 0), (1000, 256, 0), (1000, 256, 128), (2500, 512, 0), (19000, 256, 0)])
def test_f16x3_fused_gate_projection_forward(E, Hp, tile_n):
    rng = np.random.default_rng(E + Hp)
    h = torch.rand(E, Hp, device=DEV) * 2 - 1            # |h| <= 1 like o * tanh(c)
    h[0] = 0
    U4 = torch.randn(4 * Hp, Hp, device=DEV) / Hp ** 0.5
    idx, csr = _rand_csr(rng, E, E)
    planes, scal = Fn.f16_prepare_weights(U4)
    out = torch.empty(E, 4 * Hp, device=DEV)
    Fn.gemm_f16x3(h, Hp, planes, scal, out, E, 4 * Hp, Hp, a_csr=csr, split_dim=1, split_at=3 * Hp, a_amax_const=1.0,
                  a_deg=_max_deg(csr), tile_n=tile_n)
    S = _gsum(h, idx)
    ref = torch.cat([S @ U4[:3 * Hp].double().t(), h.double() @ U4[3 * Hp:].double().t()], 1)
    assert rel(out, ref) <= 2e-6, rel(out, ref)


@pytest.mark.parametrize("E,Hp,gscale", [(300, 128, 1.0), (1000, 256, 1e-7), (5000, 256, 1.0), (19000, 256, 3e3)])
def test_f16x3_fused_state_gradient(E, Hp, gscale):
    rng = np.random.default_rng(E)
    dA = torch.randn(E, 4 * Hp, device=DEV) * gscale
    dP = torch.randn(E, Hp, device=DEV) * gscale
    U4 = torch.randn(4 * Hp, Hp, device=DEV) / Hp ** 0.5
    idx, csr = _rand_csr(rng, E, E)
    bt = csr.T
    planes, scal = Fn.f16_prepare_weights(U4, transpose=True)        # [Hp, 4Hp]
    amax = torch.zeros(16, device=DEV)
    amax[3] = max(float(dA.abs().max()), float(dP.abs().max()) / max(1, int(_max_deg(bt))))
    out = torch.empty(E, Hp, device=DEV)
    Fn.gemm_f16x3(dA, 4 * Hp, planes, scal, out, E, Hp, 4 * Hp, a_csr=bt, split_dim=2, split_at=3 * Hp, A2=dP, lda2=Hp,
                  a_amax=amax, a_deg=_max_deg(bt))
    dS = dA[:, :3 * Hp].double() @ U4[:3 * Hp].double()
    scat = torch.zeros(E, Hp, dtype=torch.float64, device=DEV)
    flat = idx.reshape(-1)
    rows = torch.arange(E, device=DEV).repeat_interleave(idx.shape[1])
    keep = flat != 0
    scat.index_add_(0, flat[keep], dS[rows[keep]])
    ref = scat + dP.double() @ U4[3 * Hp:].double()
    assert rel(out, ref) <= 2e-6, rel(out, ref)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 256, 256), (5000, 256, 1024), (1000, 512, 200)])
def test_f16x3_planes_nt(M, N, K):
    """all operands by TMA, k-contiguous planes (the state-gradient GEMM dH = Z [Ug; Uf])"""
    torch.manual_seed(M)
    A = torch.randn(M, K, device=DEV) * 3e-3
    W = torch.randn(N, K, device=DEV) * 0.07
    pa, sa = Fn.f16_prepare_weights(A)
    pb, sb = Fn.f16_prepare_weights(W)
    out = torch.full((M, N), 3.0, device=DEV)
    Fn.gemm_f16x3_planes(pa, sa, pb, sb, out, M, N, K)
    assert rel(out, A.double() @ W.double().t()) <= 2e-6


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1024, 256, 300), (1024, 256, 19001), (2048, 512, 5000), (512, 128, 777),
                                   (768, 192, 1000)])
def test_f16x3_planes_tn_weight_gradient(M, N, K):
    """C += A^T B with the reduction index as the row of both operands (MN-major tiles by 3-D TMA boxes)"""
    torch.manual_seed(K)
    A = torch.randn(K, M, device=DEV) * 2.0
    B = torch.rand(K, N, device=DEV) * 2 - 1
    pa, sa = Fn.f16_prepare_weights(A)
    pb, sb = Fn.f16_prepare_weights(B)
    out = torch.ones(M, N, device=DEV)
    Fn.gemm_f16x3_planes(pa, sa, pb, sb, out, M, N, K, beta=1.0, tn=True)
    ref = A.double().t() @ B.double() + 1.0
    assert rel(out, ref) <= 2e-6, rel(out, ref)


@pytest.mark.parametrize("H,IN,F,nbr", [(20, 26, 6, 4), (250, 270, 250, 20), (37, 37, 37, 1), (250, 62, 38, 4)])
def test_lstm_pack_matches_torch_packing_and_autograd(H, IN, F, nbr):
    """ggpm_lstm_pack / ggpm_lstm_unpack vs the torch reshuffling they replace (bit-exact: pure data movement)"""
    from ggpm_b200 import packing, rnn as R
    torch.manual_seed(H + IN)
    cell = R.LSTM(IN, H, 3).to(DEV)
    Hp = Fn.padded_width(H)
    got = packing.pack_lstm(cell, F, Hp, nbr)
    Wx, T, Ug, Uf = packing.pack_lstm_torch(cell, F, Hp, nbr)
    want = (Wx, T, torch.cat([Ug, Uf], 0))
    cots = [torch.randn_like(w) for w in want]
    for g, w in zip(got, want):
        assert g.shape == w.shape and torch.equal(g, w)
    params = list(cell.parameters())
    g1 = torch.autograd.grad(sum((g * c).sum() for g, c in zip(got, cots)), params)
    g2 = torch.autograd.grad(sum((w * c).sum() for w, c in zip(want, cots)), params)
    for a, b in zip(g1, g2):
        assert rel(a, b) <= 1e-6


def test_split_pad_autograd():
    torch.manual_seed(3)
    W = torch.randn(250, 288, device=DEV, requires_grad=True)
    b = torch.randn(250, device=DEV, requires_grad=True)
    bias_p, w1, w2 = Fn.SplitPad.apply(W, b, 256, [(0, 38), (38, 288)], [40, 252])
    assert w1.shape == (256, 40) and w2.shape == (256, 252) and bias_p.shape == (256,)
    assert torch.equal(w1[:250, :38], W[:, :38]) and torch.equal(w2[:250, :250], W[:, 38:]) and torch.equal(bias_p[:250], b)
    assert float(w1[250:].abs().max()) == 0 and float(w1[:, 38:].abs().max()) == 0 and float(w2[:, 250:].abs().max()) == 0
    c1, c2, cb = torch.randn_like(w1), torch.randn_like(w2), torch.randn_like(bias_p)
    ((w1 * c1).sum() + (w2 * c2).sum() + (bias_p * cb).sum()).backward()
    assert torch.equal(W.grad, torch.cat([c1[:250, :38], c2[:250, :250]], 1)) and torch.equal(b.grad, cb[:250])


def test_linear_act_autograd(engine):
    g = torch.Generator().manual_seed(9)
    A1 = torch.randn(400, 38, generator=g, dtype=torch.float64, requires_grad=True)
    A2 = torch.randn(400, 250, generator=g, dtype=torch.float64, requires_grad=True)
    W = (torch.randn(250, 288, generator=g, dtype=torch.float64) * 0.1).requires_grad_(True)
    b = torch.randn(250, generator=g, dtype=torch.float64, requires_grad=True)
    cot = torch.randn(400, 250, generator=g, dtype=torch.float64)
    mask = torch.ones(400, 1, dtype=torch.float64)
    mask[0] = 0
    ref = torch.relu(torch.cat([A1, A2], 1) @ W.t() + b) * mask
    (ref * cot).sum().backward()
    a1, a2, w, bb = (t.detach().float().to(DEV).requires_grad_(True) for t in (A1, A2, W, b))
    out = Fn.linear_act([a1, a2], [w[:, :38], w[:, 38:]], bb, act=_lib.ACT_RELU, zero_row0=True)
    assert rel(out, ref) <= 1e-5
    (out * cot.float().to(DEV)).sum().backward()
    for got, want in ((a1, A1), (a2, A2), (w, W), (bb, b)):
        assert rel(got.grad, want.grad) <= 1e-5


def test_small_helpers():
    x = torch.randn(1000, 96, device=DEV)
    out = torch.ones(96, device=DEV)
    _lib.call("ggpm_colsum", x.data_ptr(), 96, 1000, 96, out.data_ptr(), 1, None)
    assert rel(out, x.double().sum(0) + 1) <= 1e-5
    k1 = torch.randint(0, 4, (1000,), device=DEV, dtype=torch.int32)
    k2 = torch.randint(4, 24, (1000,), device=DEV, dtype=torch.int32)
    dT = torch.zeros(24, 96, device=DEV)
    _lib.call("ggpm_table_wgrad", x.data_ptr(), 96, 1000, 96, k1.data_ptr(), k2.data_ptr(), dT.data_ptr(), 96, 24, None)
    ref = torch.zeros(24, 96, dtype=torch.float64).index_add_(0, k1.cpu().long(), x.cpu().double())
    ref.index_add_(0, k2.cpu().long(), x.cpu().double())
    assert rel(dT, ref) <= 1e-5


# ---------------------------------------------------------------------------------------------------
# whole message-passing layers against the fp64 twins (random packed weights, ragged hidden width)
# ---------------------------------------------------------------------------------------------------
def _layer_inputs(H, Hp, G, batch, seed, with_k2):
    mols = synth.random_batch(seed, batch)
    tree, graph = synth.tensorize(mols)
    fnode, fmess, agraph, bgraph = (graph if with_k2 else tree)[:4]
    E1, N1 = fmess.shape[0], fnode.shape[0]
    g = torch.Generator().manual_seed(seed)
    F = 38 if with_k2 else H
    ntab = 24 if with_k2 else 20
    colmask = torch.zeros(Hp, dtype=torch.float64)
    colmask[:H] = 1

    def gate_rows(n):          # [n, G*Hp] with zero pad columns
        return (torch.randn(n, G, Hp, generator=g, dtype=torch.float64) * 0.3 * colmask).reshape(n, G * Hp)
    phi = torch.randn(N1, F, generator=g, dtype=torch.float64)
    Wx = gate_rows(F).t().contiguous()                         # [G*Hp, F]
    T = gate_rows(ntab)
    sq = lambda: torch.randn(Hp, Hp, generator=g, dtype=torch.float64) * (0.5 / H ** 0.5) * colmask[:, None] * colmask[None]
    src = torch.from_numpy(fmess[:, 0].astype(np.int64))
    k1 = torch.from_numpy(fmess[:, 2].astype(np.int64))
    k2 = torch.from_numpy(fmess[:, 3].astype(np.int64) + 4) if with_k2 else None
    bg = torch.from_numpy(bgraph.astype(np.int64))
    cot = torch.randn(E1, Hp, generator=g, dtype=torch.float64) * colmask
    return dict(phi=phi, Wx=Wx, T=T, sq=sq, src=src, k1=k1, k2=k2, bg=bg, cot=cot, E1=E1, N1=N1, ntab=ntab)


def _message_graph(d):
    i32 = lambda t: None if t is None else t.to(torch.int32).to(DEV)
    return Fn.MessageGraph(Fn.csr_from_padded(d["bg"].to(DEV), d["E1"]), i32(d["src"]), i32(d["k1"]), i32(d["k2"]),
                           d["N1"], d["ntab"])


@pytest.mark.parametrize("H,Hp,depth,batch,with_k2", [(20, 32, 3, 3, False), (250, 256, 5, 8, True), (60, 64, 20, 4, False)])
def test_lstm_layer_forward_backward(engine, H, Hp, depth, batch, with_k2):
    d = _layer_inputs(H, Hp, 4, batch, 100 + H, with_k2)
    Ug = torch.cat([d["sq"](), d["sq"](), d["sq"]()], 0)
    Uf = d["sq"]()
    rp, col = P.csr_from_padded(d["bg"])
    Yn = d["phi"] @ d["Wx"].t()
    h_ref, hist = P.lstm_layer_fwd(Yn, d["src"], d["T"], d["k1"], d["k2"], Ug, Uf, rp, col, depth, H, Hp)
    dX, dUg, dUf = P.lstm_layer_bwd(hist, d["cot"], Ug, Uf, rp, col, depth, H, Hp)
    dYn = torch.zeros_like(Yn).index_add_(0, d["src"], dX)
    dT = torch.zeros_like(d["T"]).index_add_(0, d["k1"], dX)
    if d["k2"] is not None:
        dT.index_add_(0, d["k2"], dX)
    leaf = lambda t: t.float().to(DEV).requires_grad_(True)
    phi, Wx, T, u4 = leaf(d["phi"]), leaf(d["Wx"]), leaf(d["T"]), leaf(torch.cat([Ug, Uf], 0))
    h, c = Fn.LSTMMessages.apply(phi, Wx, T, u4, _message_graph(d), depth, H)
    assert rel(h, h_ref) <= 1e-5
    assert rel(c, hist["c"][-1]) <= 1e-5
    assert float(h[0].abs().max()) == 0 and float(h[:, H:].abs().max()) == 0
    (h * d["cot"].float().to(DEV)).sum().backward()
    assert rel(u4.grad[:3 * Hp], dUg) <= 1e-5 and rel(u4.grad[3 * Hp:], dUf) <= 1e-5
    assert rel(T.grad, dT) <= 1e-5
    assert rel(Wx.grad, dYn.t() @ d["phi"]) <= 1e-5
    assert rel(phi.grad, dYn @ d["Wx"]) <= 1e-5


@pytest.mark.parametrize("H,Hp,depth,batch,with_k2", [(20, 32, 3, 3, False), (250, 256, 4, 6, True)])
def test_gru_layer_forward_backward(engine, H, Hp, depth, batch, with_k2):
    d = _layer_inputs(H, Hp, 3, batch, 200 + H, with_k2)
    Uz, Ur, Uh = d["sq"](), d["sq"](), d["sq"]()
    br = torch.zeros(Hp, dtype=torch.float64)
    br[:H] = torch.randn(H, dtype=torch.float64, generator=torch.Generator().manual_seed(1)) * 0.1
    rp, col = P.csr_from_padded(d["bg"])
    Yn = d["phi"] @ d["Wx"].t()
    h_ref, hist = P.gru_layer_fwd(Yn, d["src"], d["T"], d["k1"], d["k2"], Uz, Ur, br, Uh, rp, col, depth, H, Hp)
    dX, dUz, dUr, dbr, dUh = P.gru_layer_bwd(hist, d["cot"], Uz, Ur, Uh, rp, col, depth, H, Hp)
    dYn = torch.zeros_like(Yn).index_add_(0, d["src"], dX)
    leaf = lambda t: t.float().to(DEV).requires_grad_(True)
    phi, Wx, T, uz, ur, b, uh = (leaf(t) for t in (d["phi"], d["Wx"], d["T"], Uz, Ur, br, Uh))
    h = Fn.GRUMessages.apply(phi, Wx, T, uz, ur, b, uh, _message_graph(d), depth, H)
    assert rel(h, h_ref) <= 1e-5
    (h * d["cot"].float().to(DEV)).sum().backward()
    for got, want in ((uz, dUz), (ur, dUr), (uh, dUh), (b, dbr), (phi, dYn @ d["Wx"]), (Wx, dYn.t() @ d["phi"])):
        assert rel(got.grad, want) <= 1e-5
