"""TEST INFRASTRUCTURE ONLY — host emulator of the C ABI in include/ggpm_b200.h.

Implements every entry point on raw HOST addresses (ctypes -> numpy -> torch views) with the torch twins
of tests/_proto.py, honouring the same buffer layouts (history slots, leading dimensions, in-place
accumulation).  tests install it through ggpm_b200._lib._install_test_backend so that the Python host
side of the product (index prep, weight packing, autograd wiring, pointer/stride arithmetic) can be
exercised on a box without a GPU.  The product never imports this file; on a GPU box the real
libggpm_b200.so is what runs (and what the -m gpu tests check against the oracle).
"""
import ctypes

import numpy as np
import torch

import _proto as P

DT = torch.float64     # emulate in fp64, store fp32


def _arr(ptr, n, ctype, npdtype):
    if n == 0:
        return np.zeros(0, dtype=npdtype)
    return np.ctypeslib.as_array((ctype * n).from_address(int(ptr)))


def _f(ptr, n):
    return torch.from_numpy(_arr(ptr, n, ctypes.c_float, np.float32))


def _i(ptr, n):
    return torch.from_numpy(_arr(ptr, n, ctypes.c_int32, np.int32))


def _tab(ptr, rows, ld, width):
    """strided [rows, width] fp32 view with leading dimension ld"""
    flat = _f(ptr, (rows - 1) * ld + width if rows > 0 else 0)
    return flat.as_strided((rows, width), (ld, 1))


def _csr(ptr_p, col_p, rows):
    rp = _i(ptr_p, rows + 1).long()
    return rp, _i(col_p, int(rp[-1])).long()


class EmuLib:
    def __init__(self):
        self.launches = 0
        self.engine = 0
        self.err = b""
        self.calls = []

    # -- library state ---------------------------------------------------------------------------
    def ggpm_last_error(self):
        return self.err

    def ggpm_abi_version(self):
        return 1

    def ggpm_set_gemm_engine(self, e):
        self.engine = e
        return 0

    def ggpm_get_gemm_engine(self):
        return self.engine

    def ggpm_launch_count(self):
        return self.launches

    def _tick(self, name):
        self.launches += 1
        self.calls.append(name)

    # -- index construction ----------------------------------------------------------------------
    def ggpm_csr_from_padded(self, idx, is64, rows, width, n_src, row_ptr, col, bad, stream):
        self._tick("csr_from_padded")
        n = rows * width
        a = _arr(idx, n, ctypes.c_int64 if is64 else ctypes.c_int32, np.int64 if is64 else np.int32)
        index = torch.from_numpy(a.astype(np.int64)).reshape(rows, width)
        oob = (index < 0) | (index >= n_src)
        if oob.any():
            _i(bad, 1)[0] |= 1
            index = index.masked_fill(oob, 0)
        rp, c = P.csr_from_padded(index)
        _i(row_ptr, rows + 1).copy_(rp.int())
        _i(col, len(c)).copy_(c.int())
        return 0

    def ggpm_check_ids(self, ids, is64, n, stride, lo, hi, bad, bit, stream):
        self._tick("check_ids")
        if n == 0:
            return 0
        a = _arr(ids, (n - 1) * stride + 1, ctypes.c_int64 if is64 else ctypes.c_int32, np.int64 if is64 else np.int32)[::stride]
        if ((a < lo) | (a >= hi)).any():
            _i(bad, 1)[0] |= bit
        return 0

    def ggpm_csr_transpose(self, row_ptr, col, rows, n_src, t_ptr, t_row, t_pos, cursor, stream):
        self._tick("csr_transpose")
        rp, c = _csr(row_ptr, col, rows)
        dest = P.csr_rows(rp)
        pos = torch.arange(len(c))
        order = torch.from_numpy(np.lexsort((pos.numpy(), c.numpy())))
        tp = torch.zeros(n_src + 1, dtype=torch.long)
        tp[1:] = torch.cumsum(torch.bincount(c, minlength=n_src), 0)
        _i(t_ptr, n_src + 1).copy_(tp.int())
        _i(t_row, len(c)).copy_(dest[order].int())
        _i(t_pos, len(c)).copy_(pos[order].int())
        return 0

    # -- primitives ------------------------------------------------------------------------------
    def ggpm_gather_sum(self, table, ld_t, row_ptr, col, out, ld_o, rows, width, accumulate, stream):
        self._tick("gather_sum")
        assert width % 4 == 0 and ld_t % 4 == 0 and ld_o % 4 == 0 and table % 16 == 0 and out % 16 == 0
        rp, c = _csr(row_ptr, col, rows)
        n_src = int(c.max()) + 1 if len(c) else 1
        t = _tab(table, n_src, ld_t, width).to(DT)
        res = P.gather_sum(t, rp, c)
        o = _tab(out, rows, ld_o, width)
        o.copy_((o.to(DT) + res if accumulate else res).float())
        return 0

    def ggpm_gemm(self, A, sam, sak, B, sbk, sbn, C, ldc, M, N, K, beta, bias, act, zero_row0, split_k, stream):
        self._tick("gemm")
        assert sam == 1 or sak == 1, "A needs a unit stride"
        assert sbk == 1 or sbn == 1, "B needs a unit stride"
        a = _f(A, (M - 1) * sam + (K - 1) * sak + 1).as_strided((M, K), (sam, sak)).to(DT)
        b = _f(B, (K - 1) * sbk + (N - 1) * sbn + 1).as_strided((K, N), (sbk, sbn)).to(DT)
        c = _tab(C, M, ldc, N)
        r = a @ b
        if beta != 0:
            r = r + beta * c.to(DT)
        if split_k > 1:
            assert bias is None and act == 0 and not zero_row0
        if bias is not None:
            r = r + _f(bias, N).to(DT)
        r = {0: lambda x: x, 1: torch.relu, 2: torch.tanh, 3: torch.sigmoid}[act](r)
        if zero_row0:
            r[0] = 0
        c.copy_(r.float())
        return 0

    # -- weight packing ----------------------------------------------------------------------------
    @staticmethod
    def _ptrs(arr, n):
        return [int(arr[i]) if arr[i] else 0 for i in range(n)]

    def ggpm_lstm_pack(self, W4, b4, H, Hp, F, Fp, IN, ntab, nbr, Wx, T, U4, stream):
        self._tick("lstm_pack")
        ws = [_tab(p, H, IN + H, IN + H) for p in self._ptrs(W4, 4)]
        bs = [_f(p, H) for p in self._ptrs(b4, 4)]
        wx, t, u = _tab(Wx, 4 * Hp, Fp, Fp), _tab(T, ntab, 4 * Hp, 4 * Hp), _tab(U4, 4 * Hp, Hp, Hp)
        wx.zero_(); t.zero_(); u.zero_()
        for g in range(4):
            wx[g * Hp:g * Hp + H, :F] = ws[g][:, :F]
            u[g * Hp:g * Hp + H, :H] = ws[g][:, IN:]
            if IN > F:
                t[:, g * Hp:g * Hp + H] = ws[g][:, F:IN].t()
            t[:nbr, g * Hp:g * Hp + H] += bs[g]
        return 0

    def ggpm_lstm_unpack(self, dW4, db4, H, Hp, F, Fp, IN, nbr, dWx, dT, dU4, stream):
        self._tick("lstm_unpack")
        ntab = IN - F if IN > F else 1
        ws = [_tab(p, H, IN + H, IN + H) for p in self._ptrs(dW4, 4)]
        bs = [_f(p, H) for p in self._ptrs(db4, 4)]
        for g in range(4):
            ws[g].zero_(); bs[g].zero_()
            if dWx:
                ws[g][:, :F] = _tab(dWx, 4 * Hp, Fp, Fp)[g * Hp:g * Hp + H, :F]
            if dU4:
                ws[g][:, IN:] = _tab(dU4, 4 * Hp, Hp, Hp)[g * Hp:g * Hp + H, :H]
            if dT:
                t = _tab(dT, ntab, 4 * Hp, 4 * Hp)
                if IN > F:
                    ws[g][:, F:IN] = t[:, g * Hp:g * Hp + H].t()
                bs[g].copy_(t[:nbr, g * Hp:g * Hp + H].sum(0))
        return 0

    def ggpm_split_pad(self, W, ldw, N, K, Np, n, c0, c1, kp, out, bias, bias_p, stream):
        self._tick("split_pad")
        w = _tab(W, N, ldw, K)
        for i, o in enumerate(self._ptrs(out, n)):
            dst = _tab(o, Np, kp[i], kp[i])
            dst.zero_()
            dst[:N, :c1[i] - c0[i]] = w[:, c0[i]:c1[i]]
        if bias_p:
            bp = _f(bias_p, Np)
            bp.zero_()
            if bias:
                bp[:N] = _f(bias, N)
        return 0

    def ggpm_split_unpad(self, dW, N, K, n, c0, c1, kp, dpieces, dbias_p, dbias, stream):
        self._tick("split_unpad")
        w = _tab(dW, N, K, K)
        w.zero_()
        for i, o in enumerate(self._ptrs(dpieces, n)):
            if o:
                w[:, c0[i]:c1[i]] = _tab(o, N, kp[i], kp[i])[:, :c1[i] - c0[i]]
        if dbias:
            d = _f(dbias, N)
            d.zero_()
            if dbias_p:
                d.copy_(_f(dbias_p, N))
        return 0

    def ggpm_act_bwd(self, dY, ld_dy, Y, ld_y, dZ, ld_dz, rows, width, act, zero_row0, stream):
        self._tick("act_bwd")
        g, y = _tab(dY, rows, ld_dy, width).to(DT), _tab(Y, rows, ld_y, width).to(DT)
        if act == 1:
            g = g * (y > 0)
        elif act == 2:
            g = g * (1 - y * y)
        elif act == 3:
            g = g * y * (1 - y)
        if zero_row0:
            g[0] = 0
        _tab(dZ, rows, ld_dz, width).copy_(g.float())
        return 0

    # -- latent head (csrc/latent.cu) ---------------------------------------------------------------
    def ggpm_rsample_fwd(self, Y, ldy, var_off, eps, ld_eps, B, L, z, ldz, kl, row_kl, stream):
        self._tick("rsample_fwd")
        y = _tab(Y, B, ldy, var_off + L).to(DT)
        m, lv = y[:, :L], -y[:, var_off:var_off + L].abs()
        rows = (1.0 + lv - m * m - torch.exp(lv)).sum(1)
        _f(row_kl, B).copy_(rows.float())
        _f(kl, 1)[0] = float(-0.5 * rows.sum() / B)
        out = m if not eps else m + torch.exp(lv / 2) * _tab(eps, B, ld_eps, L).to(DT)
        _tab(z, B, ldz, L).copy_(out.float())
        return 0

    def ggpm_rsample_bwd(self, Y, ldy, var_off, eps, ld_eps, B, L, dz, ld_dz, dkl, dY, ld_dy, width, stream):
        self._tick("rsample_bwd")
        y = _tab(Y, B, ldy, var_off + L).to(DT)
        m, v = y[:, :L], y[:, var_off:var_off + L]
        lv = -v.abs()
        gk = float(_f(dkl, 1)[0]) / B if dkl else 0.0
        g = _tab(dz, B, ld_dz, L).to(DT) if dz else torch.zeros(B, L, dtype=DT)
        dlv = -0.5 * gk * (1.0 - torch.exp(lv))
        if eps and dz:
            dlv = dlv + g * _tab(eps, B, ld_eps, L).to(DT) * 0.5 * torch.exp(lv / 2)
        out = torch.zeros(B, width, dtype=DT)
        out[:, :L] = g + gk * m
        out[:, var_off:var_off + L] = -torch.sign(v) * dlv
        _tab(dY, B, ld_dy, width).copy_(out.float())
        return 0

    # -- property heads (csrc/heads.cu) -------------------------------------------------------------------------
    @staticmethod
    def _head(ptrs, h, ldx, ld_mult, B, In, Hd):
        q = [ptrs[h * 17 + i] for i in range(17)]
        d = dict(x=_tab(q[0], B, ldx, In), W1=_tab(q[1], Hd, In, In), b1=_f(q[2], Hd), w2=_f(q[3], Hd), b2=_f(q[4], 1),
                 mult=None if not q[5] else _tab(q[5], B, ld_mult, Hd), target=None if not q[6] else _f(q[6], B),
                 hid=_tab(q[7], B, Hd, Hd), out=_f(q[8], B), sq=None if not q[9] else _f(q[9], B),
                 loss=None if not q[10] else _f(q[10], 1), dout=None if not q[11] else _f(q[11], B),
                 dloss=None if not q[12] else _f(q[12], 1), q=q)
        return d

    def ggpm_mlp_heads_fwd(self, ptrs, ldx0, ldx1, ld_mult, B, In, Hd, stream):
        self._tick("mlp_heads_fwd")
        for h, ldx in ((0, ldx0), (1, ldx1)):
            d = self._head(ptrs, h, ldx, ld_mult, B, In, Hd)
            a = torch.relu(d["x"].to(DT) @ d["W1"].to(DT).t() + d["b1"].to(DT))
            d["hid"].copy_(a.float())
            am = a if d["mult"] is None else a * d["mult"].to(DT)
            y = am @ d["w2"].to(DT) + d["b2"].to(DT)
            d["out"].copy_(y.float())
            if d["target"] is not None:
                d["loss"].copy_(((y - d["target"].to(DT)) ** 2).mean().reshape(1).float())
        return 0

    def ggpm_mlp_heads_bwd(self, ptrs, db2, ldx0, ldx1, ld_mult, ld_dx, B, In, Hd, stream):
        self._tick("mlp_heads_bwd")
        for h, ldx in ((0, ldx0), (1, ldx1)):
            d = self._head(ptrs, h, ldx, ld_mult, B, In, Hd)
            q = d["q"]
            dy = torch.zeros(B, dtype=DT)
            if d["dout"] is not None:
                dy += d["dout"].to(DT)
            if d["dloss"] is not None and d["target"] is not None:
                dy += d["dloss"].to(DT) * 2.0 * (d["out"].to(DT) - d["target"].to(DT)) / B
            a = d["hid"].to(DT)
            m = torch.ones_like(a) if d["mult"] is None else d["mult"].to(DT)
            da = dy[:, None] * d["w2"].to(DT)[None, :] * m * (a > 0)
            _tab(q[13], B, ld_dx, In).copy_((da @ d["W1"].to(DT)).float())
            _tab(q[14], Hd, In, In).copy_((da.t() @ d["x"].to(DT)).float())
            _f(q[15], Hd).copy_(da.sum(0).float())
            _f(q[16], Hd).copy_(((a * m) * dy[:, None]).sum(0).float())
            _f(db2, 2)[h] = float(dy.sum())
        return 0

    def ggpm_embedding_bwd(self, dOut, ld, idx, is64, idx_stride, rows, width, n_tab, dW, ldw, stream):
        self._tick("embedding_bwd")
        n = (rows - 1) * idx_stride + 1 if rows > 0 else 0
        a = _arr(idx, n, ctypes.c_int64 if is64 else ctypes.c_int32, np.int64 if is64 else np.int32)
        ids = torch.from_numpy(a.astype(np.int64))[::idx_stride]
        w = _tab(dW, n_tab, ldw, width)
        acc = torch.zeros(n_tab, width, dtype=DT)
        if rows > 0:
            acc.index_add_(0, ids, _tab(dOut, rows, ld, width).to(DT))
        w.copy_(acc.float())
        return 0

    def ggpm_bias_act(self, C, ldc, rows, width, bias, act, zero_row0, stream):
        self._tick("bias_act")
        assert width % 4 == 0 and ldc % 4 == 0 and C % 16 == 0
        c = _tab(C, rows, ldc, width)
        x = c.to(DT)
        if bias:
            x = x + _f(bias, width).to(DT)
        x = torch.relu(x) if act == 1 else torch.tanh(x) if act == 2 else torch.sigmoid(x) if act == 3 else x
        if zero_row0:
            x[0] = 0
        c.copy_(x.float())
        return 0

    def ggpm_colsum(self, X, ld_x, rows, width, out, accumulate, stream):
        self._tick("colsum")
        s = _tab(X, rows, ld_x, width).to(DT).sum(0)
        o = _f(out, width)
        o.copy_((o.to(DT) + s if accumulate else s).float())
        return 0

    def ggpm_table_wgrad(self, dX, ld_dx, rows, width, k1, k2, dT, ld_dt, ntab, stream):
        self._tick("table_wgrad")
        assert ntab <= 32
        x = _tab(dX, rows, ld_dx, width).to(DT)
        t = _tab(dT, ntab, ld_dt, width)
        acc = t.to(DT)
        acc.index_add_(0, _i(k1, rows).long(), x)
        if k2:
            acc.index_add_(0, _i(k2, rows).long(), x)
        t.copy_(acc.float())
        return 0

    # -- LSTM ------------------------------------------------------------------------------------
    @staticmethod
    def _xtab(Yn, src, T, k1, k2, E1, G, Hp):
        s, a = _i(src, E1).long(), _i(k1, E1).long()
        b = _i(k2, E1).long() if k2 else None
        n1 = int(s.max()) + 1
        ntab = max(int(a.max()), int(b.max()) if b is not None else 0) + 1
        return P.xproj(_tab(Yn, n1, G * Hp, G * Hp).to(DT), s, _tab(T, ntab, G * Hp, G * Hp).to(DT), a, b)

    def ggpm_lstm_plan(self, Hp, depth):
        return 0

    def ggpm_lstm_fwd(self, E1, H, Hp, depth, b_ptr, b_col, Yn, src, T, k1, k2, Ug, Uf, hist_h, hist_c, hist_gp,
                      s_scratch, w_split, hist_h16, keep, h_rotate, stream):
        self._tick("lstm_fwd")
        assert Hp % 32 == 0 and Hp >= H
        rp, col = _csr(b_ptr, b_col, E1)
        X = self._xtab(Yn, src, T, k1, k2, E1, 4, Hp)
        ug, uf = _tab(Ug, 3 * Hp, Hp, Hp).to(DT), _tab(Uf, Hp, Hp, Hp).to(DT)
        slots = depth if keep else min(depth, 2)
        hh, hc = _f(hist_h, slots * E1 * Hp).view(slots, E1, Hp), _f(hist_c, slots * E1 * Hp).view(slots, E1, Hp)
        gp = _f(hist_gp, slots * E1 * 4 * Hp).view(slots, E1, 4 * Hp)
        h = c = None
        for t in range(1, depth + 1):
            cur = (t - 1) if keep else ((t - 1) & 1)
            if t > 1:
                G = P.gather_sum(h, rp, col) @ ug.t()
                Pm = h @ uf.t()
                gp[cur].copy_(torch.cat([G, Pm], 1).float())
            else:
                G = Pm = None
            h, c = P.lstm_cell_fwd(X, G, Pm, c, rp, col, H, Hp)
            hh[cur].copy_(h.float())
            hc[cur].copy_(c.float())
            h, c = hh[cur].to(DT), hc[cur].to(DT)      # carry fp32-rounded state like the device does
        return 0

    def ggpm_lstm_bwd(self, E1, H, Hp, depth, b_ptr, b_col, bt_ptr, bt_row, Yn, src, T, k1, k2, Ug, Uf, hist_h,
                      hist_c, hist_gp, dH_final, dC_final, dX, dUg, dUf, dA, dC, dP, dS, S, dHn, w_split, hist_h16, z_planes,
                      h_rotate, bt_pos, pair_buf, pair_rows, stream):
        self._tick("lstm_bwd")
        rp, col = _csr(b_ptr, b_col, E1)
        trp, tcol = _csr(bt_ptr, bt_row, E1)
        # the reverse CSR must be the transpose of the forward one
        fwd_pairs = sorted(zip(P.csr_rows(rp).tolist(), col.tolist()))
        rev_pairs = sorted(zip(tcol.tolist(), P.csr_rows(trp).tolist()))
        assert fwd_pairs == rev_pairs, "bt is not the transpose of b"
        X = self._xtab(Yn, src, T, k1, k2, E1, 4, Hp)
        ug, uf = _tab(Ug, 3 * Hp, Hp, Hp).to(DT), _tab(Uf, Hp, Hp, Hp).to(DT)
        hh = _f(hist_h, depth * E1 * Hp).view(depth, E1, Hp).to(DT)
        hc = _f(hist_c, depth * E1 * Hp).view(depth, E1, Hp).to(DT)
        gp = _f(hist_gp, depth * E1 * 4 * Hp).view(depth, E1, 4 * Hp).to(DT)
        hist = dict(X=X, h=list(hh), c=list(hc), G=[None] + [gp[t][:, :3 * Hp] for t in range(1, depth)],
                    P=[None] + [gp[t][:, 3 * Hp:] for t in range(1, depth)])
        dx, dug, duf = P.lstm_layer_bwd(hist, _tab(dH_final, E1, Hp, Hp).to(DT), ug, uf, rp, col, depth, H, Hp,
                                        dC_final=None if not dC_final else _tab(dC_final, E1, Hp, Hp).to(DT))
        _tab(dX, E1, 4 * Hp, 4 * Hp).copy_(dx.float())
        g = _tab(dUg, 3 * Hp, Hp, Hp)
        g.copy_((g.to(DT) + dug).float())
        f = _tab(dUf, Hp, Hp, Hp)
        f.copy_((f.to(DT) + duf).float())
        return 0

    # -- GRU -------------------------------------------------------------------------------------
    def ggpm_gru_fwd(self, E1, H, Hp, depth, b_ptr, b_col, Yn, src, T, k1, k2, Uz, Ur, br, Uh, hist_h, hist_aux,
                     keep, stream):
        self._tick("gru_fwd")
        rp, col = _csr(b_ptr, b_col, E1)
        X = self._xtab(Yn, src, T, k1, k2, E1, 3, Hp)
        uz, ur, uh = (_tab(p, Hp, Hp, Hp).to(DT) for p in (Uz, Ur, Uh))
        b = _f(br, Hp).to(DT)
        # run the twin with an already-projected X: feed it through identity tables
        ident_src = torch.arange(E1)
        h, hist = P.gru_layer_fwd(X, ident_src, torch.zeros(1, 3 * Hp, dtype=DT), torch.zeros(E1, dtype=torch.long),
                                  None, uz, ur, b, uh, rp, col, depth, H, Hp)
        nslots = depth + 1 if keep else 2
        hh = _f(hist_h, nslots * E1 * Hp).view(nslots, E1, Hp)
        aux = _f(hist_aux, (depth if keep else 1) * E1 * 5 * Hp).view(-1, E1, 5 * Hp)
        if keep:
            for t in range(depth):
                hh[t].copy_(hist["h"][t].float())
                aux[t].copy_(torch.cat([hist[k][t] for k in ("S", "Gz", "Gh", "Pr", "Sg")], 1).float())
            hh[depth].copy_(h.float())
        else:
            hh[depth & 1].copy_(h.float())
        return 0

    def ggpm_gru_bwd(self, E1, H, Hp, depth, b_ptr, b_col, Yn, src, T, k1, k2, Uz, Ur, Uh, hist_h, hist_aux,
                     dH_final, dX, dUz, dUr, dbr, dUh, dA, dS, dSg, dPr, dHn, stream):
        self._tick("gru_bwd")
        rp, col = _csr(b_ptr, b_col, E1)
        X = self._xtab(Yn, src, T, k1, k2, E1, 3, Hp)
        uz, ur, uh = (_tab(p, Hp, Hp, Hp).to(DT) for p in (Uz, Ur, Uh))
        hh = _f(hist_h, (depth + 1) * E1 * Hp).view(depth + 1, E1, Hp).to(DT)
        aux = _f(hist_aux, depth * E1 * 5 * Hp).view(depth, E1, 5, Hp).to(DT)
        hist = dict(X=X, h=list(hh[:depth]))
        for i, k in enumerate(("S", "Gz", "Gh", "Pr", "Sg")):
            hist[k] = [aux[t][:, i] for t in range(depth)]
        dx, duz, dur, dbr_, duh = P.gru_layer_bwd(hist, _tab(dH_final, E1, Hp, Hp).to(DT), uz, ur, uh, rp, col,
                                                  depth, H, Hp)
        _tab(dX, E1, 3 * Hp, 3 * Hp).copy_(dx.float())
        for ptr, val in ((dUz, duz), (dUr, dur), (dUh, duh)):
            t = _tab(ptr, Hp, Hp, Hp)
            t.copy_((t.to(DT) + val).float())
        b = _f(dbr, Hp)
        b.copy_((b.to(DT) + dbr_).float())
        return 0

    # -- incremental message passing (sparse.cu) ---------------------------------------------------
    def ggpm_rows_gather(self, table, ld_t, idx, M, K, width, pad_to, pairs, ld_p, total, ld_s, stream):
        self._tick("rows_gather")
        assert pad_to >= width and (not pairs or ld_p >= pad_to) and (not total or ld_s >= pad_to)
        ix = _i(idx, M * K).long()
        t = _tab(table, int(ix.max()) + 1, ld_t, width)
        picked = torch.nn.functional.pad(t[ix], (0, pad_to - width))
        if pairs:
            _tab(pairs, M * K, ld_p, pad_to).copy_(picked)
        if total:
            _tab(total, M, ld_s, pad_to).copy_(picked.to(DT).reshape(M, K, pad_to).sum(1).float())
        return 0

    def ggpm_rows_scatter(self, table, ld_t, idx, M, width, src, ld_s, stream):
        self._tick("rows_scatter")
        ix = _i(idx, M).long()
        assert len(set(ix.tolist())) == M, "rows_scatter: repeated row ids"
        t = _tab(table, int(ix.max()) + 1, ld_t, width)
        t[ix] = _tab(src, M, ld_s, width) if src else torch.zeros(M, width)
        return 0

    def ggpm_rows_scatter_add(self, table, ld_t, idx, M, K, width, src, ld_s, bcast, ld_b, stream):
        self._tick("rows_scatter_add")
        ix = _i(idx, M * K).long()
        t = _tab(table, int(ix.max()) + 1, ld_t, width)
        v = torch.zeros(M * K, width, dtype=DT)
        if src:
            v = v + _tab(src, M * K, ld_s, width).to(DT)
        if bcast:
            v = v + _tab(bcast, M, ld_b, width).to(DT).repeat_interleave(K, 0)
        acc = t.to(DT)
        acc.index_add_(0, ix, v)
        t.copy_(acc.float())
        return 0

    def ggpm_sparse_lstm_cell_fwd(self, M, K, H, A, lda, Pf, Cn, h_sub, c_sub, gates, fs, stream):
        self._tick("sparse_lstm_cell_fwd")
        a = _tab(A, M, lda, 4 * H).to(DT)
        i, o, u, xf = torch.sigmoid(a[:, :H]), torch.sigmoid(a[:, H:2 * H]), torch.tanh(a[:, 2 * H:3 * H]), a[:, 3 * H:]
        f = torch.sigmoid(xf[:, None, :] + _tab(Pf, M * K, H, H).to(DT).reshape(M, K, H))
        c = i * u + (f * _tab(Cn, M * K, H, H).to(DT).reshape(M, K, H)).sum(1)
        tc = torch.tanh(c)
        _tab(h_sub, M, H, H).copy_((o * tc).float())
        _tab(c_sub, M, H, H).copy_(c.float())
        _tab(gates, M, 4 * H, 4 * H).copy_(torch.cat([i, o, u, tc], 1).float())
        _tab(fs, M * K, H, H).copy_(f.reshape(M * K, H).float())
        return 0

    def ggpm_sparse_lstm_cell_bwd(self, M, K, H, dh, dc, gates, fs, Cn, dA, dPf, dCn, dXacc, stream):
        self._tick("sparse_lstm_cell_bwd")
        g = _tab(gates, M, 4 * H, 4 * H).to(DT)
        i, o, u, tc = g[:, :H], g[:, H:2 * H], g[:, 2 * H:3 * H], g[:, 3 * H:]
        gh = _tab(dh, M, H, H).to(DT)
        dct = _tab(dc, M, H, H).to(DT) + gh * o * (1 - tc * tc)
        f = _tab(fs, M * K, H, H).to(DT).reshape(M, K, H)
        cn = _tab(Cn, M * K, H, H).to(DT).reshape(M, K, H)
        daf = dct[:, None, :] * cn * f * (1 - f)
        d = torch.cat([dct * u * i * (1 - i), gh * tc * o * (1 - o), dct * i * (1 - u * u), daf.sum(1)], 1)
        _tab(dA, M, 4 * H, 4 * H).copy_(d.float())
        _tab(dPf, M * K, H, H).copy_(daf.reshape(M * K, H).float())
        _tab(dCn, M * K, H, H).copy_((dct[:, None, :] * f).reshape(M * K, H).float())
        x = _tab(dXacc, M, 4 * H, 4 * H)
        x.copy_((x.to(DT) + d).float())
        return 0

    def ggpm_sparse_gru_gate_fwd(self, M, K, H, Xr, ldx, Pr, Hn, rs, Sg, stream):
        self._tick("sparse_gru_gate_fwd")
        r = torch.sigmoid(_tab(Xr, M, ldx, H).to(DT)[:, None, :] + _tab(Pr, M * K, H, H).to(DT).reshape(M, K, H))
        sg = (r * _tab(Hn, M * K, H, H).to(DT).reshape(M, K, H)).sum(1)
        _tab(rs, M * K, H, H).copy_(r.reshape(M * K, H).float())
        _tab(Sg, M, H, H).copy_(sg.float())
        return 0

    def ggpm_sparse_gru_gate_bwd(self, M, K, H, dSg, rs, Hn, dPr, dHn, dXr, ldx, stream):
        self._tick("sparse_gru_gate_bwd")
        g = _tab(dSg, M, H, H).to(DT)[:, None, :]
        r = _tab(rs, M * K, H, H).to(DT).reshape(M, K, H)
        hn = _tab(Hn, M * K, H, H).to(DT).reshape(M, K, H)
        dar = g * hn * r * (1 - r)
        _tab(dPr, M * K, H, H).copy_(dar.reshape(M * K, H).float())
        _tab(dHn, M * K, H, H).copy_((g * r).reshape(M * K, H).float())
        x = _tab(dXr, M, ldx, H)
        x.copy_((x.to(DT) + dar.sum(1)).float())
        return 0

    def ggpm_sparse_gru_out_fwd(self, M, H, A, lda, S, h_sub, zh, stream):
        self._tick("sparse_gru_out_fwd")
        a = _tab(A, M, lda, 2 * H).to(DT)
        z, ht = torch.sigmoid(a[:, :H]), torch.tanh(a[:, H:])
        s = _tab(S, M, H, H).to(DT)
        _tab(h_sub, M, H, H).copy_(((1 - z) * s + z * ht).float())
        _tab(zh, M, 2 * H, 2 * H).copy_(torch.cat([z, ht], 1).float())
        return 0

    def ggpm_sparse_gru_out_bwd(self, M, H, dh, zh, S, dA, dS, dXacc, ldx, stream):
        self._tick("sparse_gru_out_bwd")
        zz = _tab(zh, M, 2 * H, 2 * H).to(DT)
        z, ht = zz[:, :H], zz[:, H:]
        g = _tab(dh, M, H, H).to(DT)
        d = torch.cat([g * (ht - _tab(S, M, H, H).to(DT)) * z * (1 - z), g * z * (1 - ht * ht)], 1)
        _tab(dA, M, 2 * H, 2 * H).copy_(d.float())
        _tab(dS, M, H, H).copy_((g * (1 - z)).float())
        x = _tab(dXacc, M, ldx, 2 * H)
        x.copy_((x.to(DT) + d).float())
        return 0

    # -- whole-call drivers (mirror the sequencing of sparse.cu on host addresses) ------------------
    def _g(self, A, lda, W, ldw, C, ldc, M, N, K, beta):          # C (+)= A @ W^T
        return self.ggpm_gemm(A, lda, 1, W, 1, ldw, C, ldc, M, N, K, beta, None, 0, 0, 1, None)

    def _gb(self, dY, ldy, W, ldw, dX, ldx, M, N, K, beta):       # dX (+)= dY @ W
        return self.ggpm_gemm(dY, ldy, 1, W, ldw, 1, dX, ldx, M, N, K, beta, None, 0, 0, 1, None)

    def _gw(self, dY, ldy, X, ldx, dW, ldw, N, Kw, rows):         # dW += dY^T @ X
        return self.ggpm_gemm(dY, 1, ldy, X, ldx, 1, dW, ldw, N, Kw, rows, 1.0, None, 0, 0, 1, None)

    def ggpm_sparse_lstm_fwd(self, M, K, H, Hp, depth, hw, cw, ld_t, sub, nbr, Xg, U4, S, Hn, Cn, gates, fs, A, h_sub,
                             c_sub, stream):
        row, pair = 4 * M * Hp, 4 * M * K * Hp          # bytes
        self.ggpm_rows_scatter(hw, ld_t, sub, M, H, None, 0, None)
        self.ggpm_rows_scatter(cw, ld_t, sub, M, H, None, 0, None)
        for t in range(depth):
            S_t, Hn_t, Cn_t, fs_t = S + t * row, Hn + t * pair, Cn + t * pair, fs + t * pair
            self.ggpm_rows_gather(hw, ld_t, nbr, M, K, H, Hp, Hn_t, Hp, S_t, Hp, None)
            self.ggpm_rows_gather(cw, ld_t, nbr, M, K, H, Hp, Cn_t, Hp, None, 0, None)
            _f(A, 4 * M * Hp).copy_(_f(Xg, 4 * M * Hp))
            self._g(S_t, Hp, U4, Hp, A, 4 * Hp, M, 3 * Hp, Hp, 1.0)
            self._g(Hn_t, Hp, U4 + 4 * 3 * Hp * Hp, Hp, fs_t, Hp, M * K, Hp, Hp, 0.0)
            self.ggpm_sparse_lstm_cell_fwd(M, K, Hp, A, 4 * Hp, fs_t, Cn_t, h_sub, c_sub, gates + t * 4 * row, fs_t, None)
            self.ggpm_rows_scatter(hw, ld_t, sub, M, H, h_sub, Hp, None)
            self.ggpm_rows_scatter(cw, ld_t, sub, M, H, c_sub, Hp, None)
        return 0

    def ggpm_sparse_lstm_bwd(self, M, K, H, Hp, depth, dHw, dCw, ld_t, sub, nbr, U4, S, Hn, Cn, gates, fs, dXg, dU4, dA,
                             dPf, dCn, dh_sub, dc_sub, dS, dHn, stream):
        row, pair = 4 * M * Hp, 4 * M * K * Hp
        Uf, dUf = U4 + 4 * 3 * Hp * Hp, dU4 + 4 * 3 * Hp * Hp
        _f(dXg, 4 * M * Hp).zero_()
        _f(dU4, 4 * Hp * Hp).zero_()
        for t in reversed(range(depth)):
            S_t, Hn_t, Cn_t, fs_t = S + t * row, Hn + t * pair, Cn + t * pair, fs + t * pair
            self.ggpm_rows_gather(dHw, ld_t, sub, M, 1, H, Hp, dh_sub, Hp, None, 0, None)
            self.ggpm_rows_gather(dCw, ld_t, sub, M, 1, H, Hp, dc_sub, Hp, None, 0, None)
            self.ggpm_rows_scatter(dHw, ld_t, sub, M, H, None, 0, None)
            self.ggpm_rows_scatter(dCw, ld_t, sub, M, H, None, 0, None)
            self.ggpm_sparse_lstm_cell_bwd(M, K, Hp, dh_sub, dc_sub, gates + t * 4 * row, fs_t, Cn_t, dA, dPf, dCn, dXg, None)
            self._gb(dA, 4 * Hp, U4, Hp, dS, Hp, M, Hp, 3 * Hp, 0.0)
            self._gb(dPf, Hp, Uf, Hp, dHn, Hp, M * K, Hp, Hp, 0.0)
            self._gw(dA, 4 * Hp, S_t, Hp, dU4, Hp, 3 * Hp, Hp, M)
            self._gw(dPf, Hp, Hn_t, Hp, dUf, Hp, Hp, Hp, M * K)
            self.ggpm_rows_scatter_add(dHw, ld_t, nbr, M, K, H, dHn, Hp, dS, Hp, None)
            self.ggpm_rows_scatter_add(dCw, ld_t, nbr, M, K, H, dCn, Hp, None, 0, None)
        self.ggpm_rows_scatter(dHw, ld_t, sub, M, H, None, 0, None)
        self.ggpm_rows_scatter(dCw, ld_t, sub, M, H, None, 0, None)
        return 0

    def ggpm_sparse_gru_fwd(self, M, K, H, Hp, depth, hw, ld_t, sub, nbr, Xg, Uz, Ur, Uh, S, Hn, rs, Sg, zh, A, h_sub, stream):
        row, pair = 4 * M * Hp, 4 * M * K * Hp
        self.ggpm_rows_scatter(hw, ld_t, sub, M, H, None, 0, None)
        for t in range(depth):
            S_t, Hn_t, rs_t, Sg_t = S + t * row, Hn + t * pair, rs + t * pair, Sg + t * row
            self.ggpm_rows_gather(hw, ld_t, nbr, M, K, H, Hp, Hn_t, Hp, S_t, Hp, None)
            self._g(Hn_t, Hp, Ur, Hp, rs_t, Hp, M * K, Hp, Hp, 0.0)
            self.ggpm_sparse_gru_gate_fwd(M, K, Hp, Xg + 4 * 2 * Hp, 3 * Hp, rs_t, Hn_t, rs_t, Sg_t, None)
            _tab(A, M, 2 * Hp, 2 * Hp).copy_(_tab(Xg, M, 3 * Hp, 2 * Hp))
            self._g(S_t, Hp, Uz, Hp, A, 2 * Hp, M, Hp, Hp, 1.0)
            self._g(Sg_t, Hp, Uh, Hp, A + 4 * Hp, 2 * Hp, M, Hp, Hp, 1.0)
            self.ggpm_sparse_gru_out_fwd(M, Hp, A, 2 * Hp, S_t, h_sub, zh + t * 2 * row, None)
            self.ggpm_rows_scatter(hw, ld_t, sub, M, H, h_sub, Hp, None)
        return 0

    def ggpm_sparse_gru_bwd(self, M, K, H, Hp, depth, dHw, ld_t, sub, nbr, Uz, Ur, Uh, S, Hn, rs, Sg, zh, dXg, dUz, dUr, dUh,
                            dA, dS, dSg, dPr, dHn, dh_sub, stream):
        row, pair = 4 * M * Hp, 4 * M * K * Hp
        _f(dXg, 3 * M * Hp).zero_()
        for d in (dUz, dUr, dUh):
            _f(d, Hp * Hp).zero_()
        for t in reversed(range(depth)):
            S_t, Hn_t, rs_t, Sg_t = S + t * row, Hn + t * pair, rs + t * pair, Sg + t * row
            self.ggpm_rows_gather(dHw, ld_t, sub, M, 1, H, Hp, dh_sub, Hp, None, 0, None)
            self.ggpm_rows_scatter(dHw, ld_t, sub, M, H, None, 0, None)
            self.ggpm_sparse_gru_out_bwd(M, Hp, dh_sub, zh + t * 2 * row, S_t, dA, dS, dXg, 3 * Hp, None)
            self._gb(dA, 2 * Hp, Uz, Hp, dS, Hp, M, Hp, Hp, 1.0)
            self._gb(dA + 4 * Hp, 2 * Hp, Uh, Hp, dSg, Hp, M, Hp, Hp, 0.0)
            self._gw(dA, 2 * Hp, S_t, Hp, dUz, Hp, Hp, Hp, M)
            self._gw(dA + 4 * Hp, 2 * Hp, Sg_t, Hp, dUh, Hp, Hp, Hp, M)
            self.ggpm_sparse_gru_gate_bwd(M, K, Hp, dSg, rs_t, Hn_t, dPr, dHn, dXg + 4 * 2 * Hp, 3 * Hp, None)
            self._gb(dPr, Hp, Ur, Hp, dHn, Hp, M * K, Hp, Hp, 1.0)
            self._gw(dPr, Hp, Hn_t, Hp, dUr, Hp, Hp, Hp, M * K)
            self.ggpm_rows_scatter_add(dHw, ld_t, nbr, M, K, H, dHn, Hp, dS, Hp, None)
        self.ggpm_rows_scatter(dHw, ld_t, sub, M, H, None, 0, None)
        return 0
