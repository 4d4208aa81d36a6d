"""TEST INFRASTRUCTURE ONLY — torch twins of the CUDA kernels, in the CUDA path's own decomposition.

The CUDA library does not run the reference's padded formulation: it hoists the x-projection out of
the depth loop, factors the per-neighbour forget/reset projection per *source* message, walks CSR
neighbour lists and hand-writes the backward pass.  This module restates that
decomposition kernel by kernel in plain torch (any dtype) so that
  * tests/test_proto_math.py proves, on CPU in fp64, that the decomposition (forward AND the manual
    backward) equals the oracle / autograd to ~1e-13, and
  * the GPU tests can compare each CUDA kernel with its twin on identical inputs.
Nothing in ggpm_b200/ imports this file.
"""
import torch


def csr_from_padded(index):
    """index: [M, K] long -> (row_ptr [M+1] long, col [nnz] long); non-zero entries in slot order."""
    keep = index != 0
    row_ptr = torch.zeros(index.shape[0] + 1, dtype=torch.long)
    row_ptr[1:] = torch.cumsum(keep.sum(1), 0)
    return row_ptr, index[keep]


def csr_rows(row_ptr):
    """destination row of every CSR entry"""
    return torch.repeat_interleave(torch.arange(len(row_ptr) - 1), row_ptr[1:] - row_ptr[:-1])


def gather_sum(table, row_ptr, col):
    out = torch.zeros(len(row_ptr) - 1, table.shape[1], dtype=table.dtype)
    out.index_add_(0, csr_rows(row_ptr), table[col])
    return out


def scatter_rows(vals, row_ptr, col, n_src):
    """transpose of gather_sum: out[n] = sum over entries (e, n) of vals[e]"""
    out = torch.zeros(n_src, vals.shape[1], dtype=vals.dtype)
    out.index_add_(0, col, vals[csr_rows(row_ptr)])
    return out


def xproj(Yn, src, T, k1, k2):
    x = Yn[src] + T[k1]
    if k2 is not None:
        x = x + T[k2]
    return x


def _mask(E1, H, Hp, dtype):
    m = torch.ones(E1, Hp, dtype=dtype)
    m[0] = 0
    m[:, H:] = 0
    return m


# ------------------------------------------------------------------------------------------------
# LSTM  (gate packing along the 4*Hp axis: [i | o | u | f])
# ------------------------------------------------------------------------------------------------

def lstm_cell_fwd(X, G, P_prev, c_prev, row_ptr, col, H, Hp):
    """One round.  G: [E1, 3Hp] or None (first round), P_prev/c_prev: [E1, Hp] or None."""
    E1 = X.shape[0]
    Xi, Xo, Xu, Xf = X[:, :Hp], X[:, Hp:2 * Hp], X[:, 2 * Hp:3 * Hp], X[:, 3 * Hp:]
    if G is None:
        gi = go = gu = 0
    else:
        gi, go, gu = G[:, :Hp], G[:, Hp:2 * Hp], G[:, 2 * Hp:]
    i, o, u = torch.sigmoid(Xi + gi), torch.sigmoid(Xo + go), torch.tanh(Xu + gu)
    chat = i * u
    if P_prev is not None:
        rows = csr_rows(row_ptr)
        f = torch.sigmoid(Xf[rows] + P_prev[col])
        chat = chat.index_add(0, rows, f * c_prev[col])
    m = _mask(E1, H, Hp, X.dtype)
    return o * torch.tanh(chat) * m, chat * m


def lstm_layer_fwd(Yn, src, T, k1, k2, Ug, Uf, row_ptr, col, depth, H, Hp):
    """Returns final h and the history (h_t, c_t, G_t, P_t lists, 1-based round t at index t-1)."""
    X = xproj(Yn, src, T, k1, k2)
    hs, cs, Gs, Ps = [], [], [], []
    h = c = None
    for t in range(1, depth + 1):
        if t > 1:
            S = gather_sum(h, row_ptr, col)
            G = S @ Ug.t()
            P = h @ Uf.t()
        else:
            G = P = None
        h, c = lstm_cell_fwd(X, G, P, c, row_ptr, col, H, Hp)
        hs.append(h); cs.append(c); Gs.append(G); Ps.append(P)   # Ps[t-1] = U_f h^{t-1}
    return h, dict(X=X, h=hs, c=cs, G=Gs, P=Ps)


def lstm_cell_bwd(X, G, P_prev, c_prev, c_cur, dH, dC, row_ptr, col, H, Hp):
    """Returns dA [E1, 4Hp] = [dai | dao | dau | sum_n daf_en], dP_prev [E1,Hp], dC_prev [E1,Hp]."""
    E1 = X.shape[0]
    m = _mask(E1, H, Hp, X.dtype)
    dH, dC = dH * m, dC * m
    Xi, Xo, Xu, Xf = X[:, :Hp], X[:, Hp:2 * Hp], X[:, 2 * Hp:3 * Hp], X[:, 3 * Hp:]
    if G is None:
        gi = go = gu = 0
    else:
        gi, go, gu = G[:, :Hp], G[:, Hp:2 * Hp], G[:, 2 * Hp:]
    i, o, u = torch.sigmoid(Xi + gi), torch.sigmoid(Xo + go), torch.tanh(Xu + gu)
    th = torch.tanh(c_cur)                 # c_cur == chat on unmasked entries; masked ones carry zero grads
    dchat = dC + dH * o * (1 - th * th)
    dai = dchat * u * i * (1 - i)
    dao = dH * th * o * (1 - o)
    dau = dchat * i * (1 - u * u)
    daf_sum = torch.zeros_like(dai)
    dP = torch.zeros_like(dai)
    dCp = torch.zeros_like(dai)
    if P_prev is not None:
        rows = csr_rows(row_ptr)
        f = torch.sigmoid(Xf[rows] + P_prev[col])
        daf = dchat[rows] * c_prev[col] * f * (1 - f)
        daf_sum.index_add_(0, rows, daf)
        dP.index_add_(0, col, daf)
        dCp.index_add_(0, col, dchat[rows] * f)
    return torch.cat([dai, dao, dau, daf_sum], 1), dP, dCp


def lstm_layer_bwd(hist, dH_final, Ug, Uf, row_ptr, col, depth, H, Hp, dC_final=None):
    """Returns dX [E1, 4Hp], dUg [3Hp, Hp], dUf [Hp, Hp]."""
    X = hist["X"]
    E1 = X.shape[0]
    dH = dH_final.clone()
    dC = torch.zeros_like(dH) if dC_final is None else dC_final.clone()
    dX = torch.zeros_like(X)
    dUg = torch.zeros_like(Ug)
    dUf = torch.zeros_like(Uf)
    for t in range(depth, 0, -1):
        h_prev = hist["h"][t - 2] if t > 1 else None
        c_prev = hist["c"][t - 2] if t > 1 else None
        dA, dP, dCp = lstm_cell_bwd(X, hist["G"][t - 1], hist["P"][t - 1], c_prev, hist["c"][t - 1], dH, dC,
                                    row_ptr, col, H, Hp)
        dX += dA
        if t == 1:
            break
        S = gather_sum(h_prev, row_ptr, col)
        dUg += dA[:, :3 * Hp].t() @ S
        dUf += dP.t() @ h_prev
        dS = dA[:, :3 * Hp] @ Ug
        dH = scatter_rows(dS, row_ptr, col, E1) + dP @ Uf
        dC = dCp
    return dX, dUg, dUf


# ------------------------------------------------------------------------------------------------
# GRU  (x-gate packing along the 3*Hp axis: [z | h | r])
# ------------------------------------------------------------------------------------------------

def gru_layer_fwd(Yn, src, T, k1, k2, Uz, Ur, br, Uh, row_ptr, col, depth, H, Hp):
    X = xproj(Yn, src, T, k1, k2)
    E1 = X.shape[0]
    Xz, Xh, Xr = X[:, :Hp], X[:, Hp:2 * Hp], X[:, 2 * Hp:]
    rows = csr_rows(row_ptr)
    m = _mask(E1, H, Hp, X.dtype)
    h = torch.zeros(E1, Hp, dtype=X.dtype)
    hist = dict(X=X, h=[], S=[], Gz=[], Gh=[], Pr=[], Sg=[])
    for t in range(1, depth + 1):
        S = gather_sum(h, row_ptr, col)
        Gz = S @ Uz.t()
        Pr = h @ Ur.t() + br
        r = torch.sigmoid(Xr[rows] + Pr[col])
        Sg = torch.zeros_like(h).index_add_(0, rows, r * h[col])
        Gh = Sg @ Uh.t()
        z = torch.sigmoid(Xz + Gz)
        cand = torch.tanh(Xh + Gh)
        hist["S"].append(S); hist["Gz"].append(Gz); hist["Gh"].append(Gh); hist["Pr"].append(Pr); hist["Sg"].append(Sg)
        hist["h"].append(h)                       # h^{t-1}
        h = ((1 - z) * S + z * cand) * m
    return h, hist


def gru_layer_bwd(hist, dH_final, Uz, Ur, Uh, row_ptr, col, depth, H, Hp):
    X = hist["X"]
    E1 = X.shape[0]
    Xz, Xh, Xr = X[:, :Hp], X[:, Hp:2 * Hp], X[:, 2 * Hp:]
    rows = csr_rows(row_ptr)
    m = _mask(E1, H, Hp, X.dtype)
    dH = dH_final.clone()
    dX = torch.zeros_like(X)
    dUz, dUr, dUh = torch.zeros_like(Uz), torch.zeros_like(Ur), torch.zeros_like(Uh)
    dbr = torch.zeros(Hp, dtype=X.dtype)
    for t in range(depth, 0, -1):
        S, Gz, Gh, Pr, Sg, h_prev = (hist[k][t - 1] for k in ("S", "Gz", "Gh", "Pr", "Sg", "h"))
        dH = dH * m
        z = torch.sigmoid(Xz + Gz)
        cand = torch.tanh(Xh + Gh)
        daz = dH * (cand - S) * z * (1 - z)
        dah = dH * z * (1 - cand * cand)
        dS = dH * (1 - z) + daz @ Uz
        dSg = dah @ Uh
        r = torch.sigmoid(Xr[rows] + Pr[col])
        dar = dSg[rows] * h_prev[col] * r * (1 - r)
        dXr = torch.zeros_like(dah).index_add_(0, rows, dar)
        dPr = torch.zeros_like(dah).index_add_(0, col, dar)
        dX += torch.cat([daz, dah, dXr], 1)
        dUz += daz.t() @ S
        dUh += dah.t() @ Sg
        dUr += dPr.t() @ h_prev
        dbr += dPr.sum(0)
        dH = torch.zeros_like(dah).index_add_(0, col, dS[rows] + dSg[rows] * r) + dPr @ Ur
    return dX, dUz, dUr, dbr, dUh
