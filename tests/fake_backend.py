"""TEST INFRASTRUCTURE ONLY -- a torch-CPU emulation of kernels.Backend (the C-ABI semantics of
include/odex.h), so the HOST-side logic of the product (solver loops, speculative dopri5 chunks, hand-written
backward replay, layout handling) can be checked against the oracle on a box without a GPU.  It is injected
with kernels.set_backend() by the `fake_backend` fixture only; the product never imports this file and has no
CPU path of its own.  What this cannot check -- the CUDA kernels themselves -- is covered by the `-m gpu` tests.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np
import torch
import torch.nn.functional as F

from ode_rl_b200 import _lib
from ode_rl_b200.kernels import ACT_NONE, ACT_RELU, ACT_TANH

C_ERROR = [35 / 384 - 1951 / 21600, 0.0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
           -2187 / 6784 - -12231 / 42400, 11 / 84 - 649 / 6300, -1.0 / 60.0]
C_MID = [6025192743 / 30085553152 / 2, 0.0, 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
         187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2]

f32 = np.float32


def _skipped(skip):
    return skip is not None and int(skip.reshape(-1)[0]) != 0


def _act(v, kind):
    if kind == ACT_RELU:
        return torch.relu(v)
    if kind == ACT_TANH:
        return torch.tanh(v)
    return v


def _dact(a, kind):
    if kind == ACT_RELU:
        return (a > 0).to(a.dtype)
    if kind == ACT_TANH:
        return 1 - a * a
    return torch.ones_like(a)


class _State:
    """ctypes view of the odex_dopri_state header living in a CPU uint8 tensor."""

    def __init__(self, st: torch.Tensor):
        self.arr = st.numpy()
        self.s = _lib.DopriState.from_buffer(self.arr)


class FakeBackend:
    name = "fake-cpu"

    def __init__(self):
        self.launches = 0
        self.wgrad_scaled = 0

    @staticmethod
    def require(t):
        return None

    # ------------------------------------------------------------------ weights / conv
    def pack_weights(self, w, dgrad, tc):
        p = w.detach().clone()
        p._odex_dgrad = bool(dgrad)
        return p

    def conv2d(self, x0, x1, w_simt, w_tc, N, R, bias, act, out, dact=None, dact_kind=0, res=None, alpha=1.0,
               alpha_dev=None, skip=None, impl=0):
        self.launches += 1
        if _skipped(skip):
            return out
        w = w_tc if w_tc is not None else w_simt
        assert w is not None, "conv2d: no packed weights"
        wt = w
        if w._odex_dgrad:
            wt = w.transpose(0, 1).flip(2, 3)
        x = x0 if x1 is None else torch.cat((x0, x1), 3)
        assert wt.shape[0] == N and wt.shape[1] == x.shape[3], (wt.shape, N, x.shape)
        v = F.conv2d(x.permute(0, 3, 1, 2), wt, None if bias is None else bias.detach(), padding=(R - 1) // 2)
        v = _act(v, act).permute(0, 2, 3, 1)
        if dact is not None:
            v = v * _dact(dact, dact_kind)
        if res is not None:
            a = f32(alpha)
            if alpha_dev is not None:
                a = f32(a * f32(float(alpha_dev.reshape(-1)[0])))
            v = res + float(a) * v
        out.copy_(v)
        return out

    # ------------------------------------------------------------------ decoder tail
    @staticmethod
    def decode_tail_supported(p):
        return p.dim() == 4 and tuple(p.shape[1:]) == (16, 16, 128) and p.shape[0] > 0 and p.is_contiguous()

    @staticmethod
    def _tail(p, w, b, slope):
        n, hc, wc, c4 = p.shape
        c = c4 // 4
        h = p.view(n, hc, wc, 2, 2, c).permute(0, 5, 1, 3, 2, 4).reshape(n, c, 2 * hc, 2 * wc)      # depth-to-space
        return F.conv_transpose2d(F.leaky_relu(h, slope), w.reshape(c, 1, 4, 4), b, stride=2, padding=1)[:, 0]

    def decode_tail_fwd(self, p, w, b, slope, out):
        self.launches += 1
        out.copy_(self._tail(p.detach(), w.detach(), None if b is None else b.detach(), slope))
        return out

    def decode_tail_bwd(self, p, dout, w, slope, dp, dw, db):
        self.launches += 1
        with torch.enable_grad():
            pp = p.detach().clone().requires_grad_(True)
            ww = w.detach().clone().requires_grad_(True)
            o = self._tail(pp, ww, None, slope)
            gp, gw = torch.autograd.grad(o, [pp, ww], dout.reshape(o.shape))
        dp.copy_(gp)
        if dw is not None:
            dw += gw.reshape(dw.shape)
        if db is not None:
            db += dout.sum()

    # ------------------------------------------------------------------ encoder stem
    @staticmethod
    def encode_stem_supported(x, cout):
        return x.dim() == 4 and tuple(x.shape[1:]) == (1, 64, 64) and cout == 16 and x.shape[0] > 0

    @staticmethod
    def _stem(x, w, b, slope):
        y = F.leaky_relu(F.conv2d(x, w.reshape(16, 1, 3, 3), b, stride=2, padding=1), slope)       # [n,16,32,32]
        n = y.shape[0]
        return y.reshape(n, 16, 16, 2, 16, 2).permute(0, 2, 4, 3, 5, 1).reshape(n, 16, 16, 64)

    def encode_stem_fwd(self, x, w, b, slope, cells):
        self.launches += 1
        cells.copy_(self._stem(x.detach(), w.detach(), None if b is None else b.detach(), slope))
        return cells

    def encode_stem_bwd(self, x, cells, dcells, slope, dw, db):
        self.launches += 1
        with torch.enable_grad():
            ww = torch.zeros(16, 9).requires_grad_(True)
            bb = torch.zeros(16).requires_grad_(True)
            # leaky' from the saved output's sign; the linear part's gradient does not depend on the weights
            mask = torch.where(cells > 0, torch.ones_like(cells), torch.full_like(cells, slope))
            n = x.shape[0]
            pre = F.conv2d(x.detach(), ww.reshape(16, 1, 3, 3), bb, stride=2, padding=1)
            pre = pre.reshape(n, 16, 16, 2, 16, 2).permute(0, 2, 4, 3, 5, 1).reshape(n, 16, 16, 64)
            gw, gb = torch.autograd.grad(pre, [ww, bb], dcells.detach() * mask)
        if dw is not None:
            dw += gw.reshape(dw.shape)
        if db is not None:
            db += gb

    # ------------------------------------------------------------------ bilinear 2x upsample
    def upsample2x_fwd(self, x, out):
        self.launches += 1
        out.copy_(F.interpolate(x.detach().permute(0, 3, 1, 2), scale_factor=2, mode="bilinear", align_corners=False).permute(0, 2, 3, 1))
        return out

    def upsample2x_bwd(self, dout, dx):
        self.launches += 1
        with torch.enable_grad():
            z = torch.zeros(dx.shape, dtype=dx.dtype).permute(0, 3, 1, 2).requires_grad_(True)
            o = F.interpolate(z, scale_factor=2, mode="bilinear", align_corners=False)
            (g,) = torch.autograd.grad(o, [z], dout.detach().permute(0, 3, 1, 2))
        dx.copy_(g.permute(0, 2, 3, 1))
        return dx

    def conv2d_wgrad(self, x0, x1, dy, R, dw, db):
        self.launches += 1
        x = x0 if x1 is None else torch.cat((x0, x1), 3)
        g = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2).contiguous(), dw.shape, dy.permute(0, 3, 1, 2).contiguous(),
                                        padding=(R - 1) // 2)
        dw += g
        if db is not None:
            db += dy.sum((0, 1, 2))

    @staticmethod
    def wgrad_f16_shape_ok(c0, c1, N, H, W, R):
        return R in (1, 3, 5) and c0 % 64 == 0 and c1 % 64 == 0 and N % 64 == 0 and W % 16 == 0 and H % 4 == 0

    # ------------------------------------------------------------------ fused conv chain (odex_convnet)
    def convnet_supported(self, x, n_layers):
        B, H, W, Cc = x.shape
        return H == 16 and W == 16 and Cc == 64 and 1 <= n_layers <= 6 and B > 0

    def convnet(self, x, w_tc, bias, act, aux, aux_kind, out, alpha=1.0, alpha_dev=None, skip=None, split=0, amax_out=None):
        self.launches += 1
        if _skipped(skip):
            return
        assert amax_out is None or split == 1
        a = x
        for l in range(len(w_tc)):
            if amax_out is not None and amax_out[l] is not None:
                amax_out[l].copy_(a.detach().abs().amax(dim=(1, 2, 3)))
            w = w_tc[l]
            wt = w.transpose(0, 1).flip(2, 3) if w._odex_dgrad else w
            v = F.conv2d(a.permute(0, 3, 1, 2), wt, None if bias[l] is None else bias[l].detach(), padding=1)
            v = _act(v, act[l]).permute(0, 2, 3, 1)
            kind = aux_kind[l] if aux[l] is not None else 0
            if kind in (1, 2):
                v = v * _dact(aux[l], kind)
            elif kind == 3:
                al = f32(alpha)
                if alpha_dev is not None:
                    al = f32(al * f32(float(alpha_dev.reshape(-1)[0])))
                v = aux[l] + float(al) * v
            v = v.contiguous()
            if out[l] is not None:
                out[l].copy_(v)
            a = v

    def conv2d_wgrad_packed(self, x0, x1, dy, R, dwp, db, impl=0, f16=False, amax=None):
        self.launches += 1
        x = x0 if x1 is None else torch.cat((x0, x1), 3)
        if amax is not None:            # the scaled kernel trusts these per-image bounds: a too-small one overflows its fp16 operands
            self.wgrad_scaled += 1
            ax, ad = amax
            assert ax.shape == (x.shape[0],) and ad.shape == (x.shape[0],), (ax.shape, ad.shape, x.shape)
            mx, md = x.abs().amax(dim=(1, 2, 3)), dy.abs().amax(dim=(1, 2, 3))
            assert bool((ax >= mx).all()) and bool((ad >= md).all()), ("wgrad maxima are not upper bounds", ax, mx, ad, md)
            assert bool((ax <= 2 * mx).all()) and bool((ad <= 2 * md).all()), ("wgrad maxima are not the operands' own", ax, mx, ad, md)
        N, Cc = dy.shape[3], x.shape[3]
        g = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2).contiguous(), (N, Cc, R, R),
                                        dy.permute(0, 3, 1, 2).contiguous(), padding=(R - 1) // 2)
        dwp += g.permute(2, 3, 0, 1).reshape(R * R, N, Cc)
        if db is not None:
            db += dy.sum((0, 1, 2))

    def wgrad_unpack(self, dwp, dw):
        self.launches += 1
        N, Cc, R, _ = dw.shape
        dw += dwp.reshape(R, R, N, Cc).permute(2, 3, 0, 1)

    # ------------------------------------------------------------------ ConvGRU gates
    @staticmethod
    def _gn(x, gamma, beta, eps):
        B, H, W, Cc = x.shape
        xg = x.reshape(B, H * W, Cc // 32, 32)
        mean = xg.mean((1, 3), keepdim=True)
        var = xg.var((1, 3), unbiased=False, keepdim=True)
        rstd = 1 / torch.sqrt(var + eps)
        y = ((xg - mean) * rstd).reshape(B, H, W, Cc) * gamma + beta
        return y, mean.reshape(B, -1), rstd.reshape(B, -1)

    def gru_gates_fwd(self, g, h, gamma, beta, eps, z, rh, stats):
        self.launches += 1
        Cc = g.shape[3] // 2
        y, mean, rstd = self._gn(g, gamma.detach(), beta.detach(), eps)
        a = torch.sigmoid(y)
        z.copy_(a[..., :Cc])
        rh.copy_(a[..., Cc:] * h)
        stats.copy_(torch.stack((mean, rstd), -1).reshape(-1))

    def gru_out_fwd(self, cpre, z, h, gamma, beta, eps, hnew, cand, stats):
        self.launches += 1
        y, mean, rstd = self._gn(cpre, gamma.detach(), beta.detach(), eps)
        c = torch.tanh(y)
        if cand is not None:
            cand.copy_(c)
        hnew.copy_((1 - z) * h + z * c)
        stats.copy_(torch.stack((mean, rstd), -1).reshape(-1))

    def gru_out_bwd(self, dhnew, cpre, cand, z, h, gamma, stats, dcpre, dz, dh, dgamma, dbeta):
        self.launches += 1
        # exact: use saved stats (mean, rstd) like the kernel does
        B, H, W, Cc = cpre.shape
        st = stats.reshape(B, Cc // 32, 2)
        mean, rstd = st[..., 0], st[..., 1]
        xg = cpre.reshape(B, H * W, Cc // 32, 32)
        xh = ((xg - mean[:, None, :, None]) * rstd[:, None, :, None]).reshape(B, H, W, Cc)
        dy = dhnew * z * (1 - cand * cand)
        dgamma += (dy * xh).sum((0, 1, 2))
        dbeta += dy.sum((0, 1, 2))
        dxh = (dy * gamma.detach()).reshape(B, H * W, Cc // 32, 32)
        xhg = xh.reshape(B, H * W, Cc // 32, 32)
        m1 = dxh.mean((1, 3), keepdim=True)
        m2 = (dxh * xhg).mean((1, 3), keepdim=True)
        dcpre.copy_((rstd[:, None, :, None] * (dxh - m1 - xhg * m2)).reshape(B, H, W, Cc))
        dz.copy_(dhnew * (cand - h))
        dh.copy_(dhnew * (1 - z))

    def gru_gates_bwd(self, dz, drh, dh_in, g, h, gamma, beta, stats, dg, dh_out, dgamma, dbeta):
        self.launches += 1
        B, H, W, C2 = g.shape
        Cc = C2 // 2
        st = stats.reshape(B, C2 // 32, 2)
        mean, rstd = st[..., 0], st[..., 1]
        xg = g.reshape(B, H * W, C2 // 32, 32)
        xh = ((xg - mean[:, None, :, None]) * rstd[:, None, :, None]).reshape(B, H, W, C2)
        a = torch.sigmoid(xh * gamma.detach() + beta.detach())
        drh_v = drh.clone()
        da = torch.cat((dz, drh_v * h), 3)
        dy = da * a * (1 - a)
        dgamma += (dy * xh).sum((0, 1, 2))
        dbeta += dy.sum((0, 1, 2))
        dxh = (dy * gamma.detach()).reshape(B, H * W, C2 // 32, 32)
        xhg = xh.reshape(B, H * W, C2 // 32, 32)
        m1 = dxh.mean((1, 3), keepdim=True)
        m2 = (dxh * xhg).mean((1, 3), keepdim=True)
        dg.copy_((rstd[:, None, :, None] * (dxh - m1 - xhg * m2)).reshape(B, H, W, C2))
        dh_out.copy_(dh_in + drh_v * a[..., Cc:])

    # ------------------------------------------------------------------ RK planes
    def lincomb(self, out, ins, coefs, dtmask=0, dt_dev=None, skip=None):
        self.launches += 1
        if _skipped(skip):
            return out
        dt = f32(float(dt_dev.reshape(-1)[0])) if dt_dev is not None else f32(1)
        acc = None
        for j, (t, c) in enumerate(zip(ins, coefs)):
            cj = f32(c) * dt if (dtmask >> j) & 1 else f32(c)
            term = t.reshape(out.shape) * float(cj)
            acc = term if acc is None else acc + term
        out.copy_(acc)
        return out

    def act_bwd(self, out, dy, a, kind):
        self.launches += 1
        out.copy_(dy * _dact(a, kind))
        return out

    # ------------------------------------------------------------------ dopri5 controller
    def dopri_state_bytes(self):
        return _lib.STATE_HEAD_BYTES + 3 * 592 * 8 + 64

    def dopri_init(self, st, t_dev, n_out, rtol, atol, max_attempts, save_cap=0):
        self.launches += 1
        st.zero_()
        s = _State(st).s
        s.t0 = s.t1 = float(t_dev[0])
        s.rtol, s.atol = rtol, atol
        s.n_out, s.jnext, s.max_attempts = n_out, 1, max_attempts
        s.save_cap, s.adv_seq = save_cap, -1
        s.done = 1 if n_out <= 1 else 0

    def dopri_initnorm(self, st, phase, y0, f0, f1, fuse_ctrl=False):
        self.launches += 1
        s = _State(st).s
        if s.done:
            return
        scale = f32(s.atol) + y0.abs() * float(f32(s.rtol))
        if phase == 0:
            s.red[0] = float(((y0 / scale) ** 2).double().sum())
            s.red[1] = float(((f0 / scale) ** 2).double().sum())
        else:
            s.red[0] = float((((f1 - f0) / scale) ** 2).double().sum())
            s.red[1] = 0.0
        s.red[2] = 0.0
        s.red[3] = float(y0.numel())
        if fuse_ctrl:
            self._initctrl(s, phase)

    def dopri_initctrl(self, st, phase):
        self.launches += 1
        s = _State(st).s
        if s.done:
            return
        self._initctrl(s, phase)

    @staticmethod
    def _begin_next(s):
        s.dt_try = s.dt
        if not (s.t1 + s.dt > s.t1):
            s.status |= 1
            s.done = 1
        if s.n_attempt >= s.max_attempts:
            s.status |= 4
            s.done = 1

    def _initctrl(self, s, phase):
        if phase == 0:
            d0 = f32(math.sqrt(s.red[0] / s.red[3]))
            d1 = f32(math.sqrt(s.red[1] / s.red[3]))
            s.d0, s.d1 = float(d0), float(d1)
            h0 = f32(1e-6) if (d0 < f32(1e-5) or d1 < f32(1e-5)) else f32(0.01) * d0 / d1
            s.h0 = float(h0)
            s.n_feval += 2
        else:
            h0, d1 = f32(s.h0), f32(s.d1)
            d2 = f32(math.sqrt(s.red[0] / s.red[3])) / h0
            s.d2 = float(d2)
            if d1 <= f32(1e-15) and d2 <= f32(1e-15):
                h1 = max(f32(1e-6), h0 * f32(1e-3))
            else:
                h1 = f32(np.power(f32(0.01) / max(d1, d2), f32(0.2)))
            s.dt = float(min(f32(100) * h0, h1))
            self._begin_next(s)

    def _errnorm(self, s, y0, y1, k7):
        dt = f32(s.dt_try)
        err = None
        for j in (0, 2, 3, 4, 5, 6):
            term = k7[j] * float(dt * f32(C_ERROR[j]))
            err = term if err is None else err + term
        tol = float(f32(s.atol)) + float(f32(s.rtol)) * torch.max(y0.abs(), y1.abs())
        r = err / tol
        s.red[0] = float((r * r).double().sum())
        s.red[1] = float((~torch.isfinite(y0)).sum())
        s.red[2] = 0.0
        s.red[3] = float(y0.numel())

    def dopri_errnorm(self, st, y0, y1, k7):
        self.launches += 1
        s = _State(st).s
        if s.done:
            return
        self._errnorm(s, y0, y1, k7)

    def dopri_errctrl(self, st, t_dev, y0, y1, k7, log, log_cap, seq):
        self.launches += 1
        s = _State(st).s
        if s.done:
            return
        self._errnorm(s, y0, y1, k7)
        self._ctrl(s, t_dev, log, log_cap, seq)

    def dopri_ctrl(self, st, t_dev, log, log_cap, seq):
        self.launches += 1
        s = _State(st).s
        if s.done:
            return
        self._ctrl(s, t_dev, log, log_cap, seq)

    def _ctrl(self, s, t_dev, log, log_cap, seq):
        dt = s.dt_try
        if s.red[1] > 0:
            s.status |= 2
            s.done = 1
            return
        ratio = float(f32(math.sqrt(s.red[0] / s.red[3])))
        s.ratio = ratio
        accept = 1 if ratio <= 1.0 else 0
        s.accept = accept
        s.n_feval += 6
        a = s.n_attempt
        if log is not None and a < log_cap:
            log[a, 0], log[a, 1], log[a, 2], log[a, 3] = s.t1, dt, ratio, accept
        s.n_attempt = a + 1
        finished = False
        if accept:
            if s.save_cap > 0 and s.n_accept >= s.save_cap:
                s.status |= 8
                s.done = 1
                return
            s.t0 = s.t1
            s.t1 = s.t1 + dt
            s.n_accept += 1
            s.dt_acc = dt
            s.adv_seq = seq
            j = s.jnext
            s.jlo = j
            while j < s.n_out and not (float(t_dev[j]) > s.t1):
                j += 1
            s.jhi = j
            s.jnext = j
            finished = j >= s.n_out
        else:
            s.n_reject += 1
            s.jlo = s.jhi = s.jnext
        if ratio == 0.0:
            factor = 10.0
        else:
            dfactor = 1.0 if ratio < 1.0 else 0.2
            factor = min(10.0, max(0.9 / ratio ** 0.2, dfactor))
        s.dt = dt * factor
        if finished:
            s.done = 1
            return
        self._begin_next(s)

    def dopri_advance(self, st, t_dev, y0, y1, k7, sol, ysave, fsave, save_cap, seq):
        self.launches += 1
        s = _State(st).s
        if s.adv_seq != seq:
            return
        dt = float(f32(s.dt_acc))
        step = s.n_accept - 1
        f0, f1 = k7[0].clone(), k7[6].clone()
        if ysave is not None and step < save_cap:
            ysave[step].copy_(y0)
            fsave[step].copy_(f0)
        if s.jhi > s.jlo:
            m = None
            for j in (0, 2, 3, 4, 5, 6):
                term = k7[j] * float(f32(dt) * f32(C_MID[j]))
                m = term if m is None else m + term
            ymid = y0 + m
            ca = 2 * dt * (f1 - f0) - 8 * (y1 + y0) + 16 * ymid
            cb = dt * (5 * f0 - 3 * f1) + 18 * y0 + 14 * y1 - 32 * ymid
            cc = dt * (f1 - 4 * f0) - 11 * y0 - 5 * y1 + 16 * ymid
            cd = dt * f0
            for j in range(s.jlo, s.jhi):
                x = float(f32((float(t_dev[j]) - s.t0) / (s.t1 - s.t0)))
                sol[j].copy_((((y0 + x * cd) + x * x * cc) + x * x * x * cb) + x * x * x * x * ca)
        y0.copy_(y1)
        k7[0].copy_(f1)

    def bn_relu_fwd(self, x, gamma, beta, eps, relu, y, stats):
        self.launches += 3
        Cc = x.shape[-1]
        r = x.reshape(-1, Cc).double()
        mean, var = r.mean(0), r.var(0, unbiased=False)
        rstd = 1.0 / torch.sqrt(var + eps)
        stats[0], stats[1], stats[2] = mean.float(), rstd.float(), var.float()
        o = (x - stats[0]) * stats[1] * gamma.detach() + beta.detach()
        y.copy_(torch.relu(o) if relu else o)

    def bn_relu_bwd(self, x, y, dy, gamma, stats, relu, dx, dgb):
        self.launches += 3
        Cc = x.shape[-1]
        g = dy * (y > 0) if relu else dy
        xh = (x - stats[0]) * stats[1]
        M = x.numel() // Cc
        dgb[0] = g.reshape(-1, Cc).sum(0)
        dgb[1] = (g * xh).reshape(-1, Cc).sum(0)
        dx.copy_(gamma.detach() * stats[1] * (g - dgb[0] / M - xh * (dgb[1] / M)))

    def warp_compose_fwd(self, prev, head, lin_x, lin_y, warped, pred):
        self.launches += 1
        b, c, h, w = prev.shape
        hd = head.movedim(-1, 1)
        grid = torch.stack(torch.broadcast_tensors(lin_x.view(1, 1, w), lin_y.view(1, h, 1)), dim=-1).expand(b, h, w, 2)
        fl = torch.cat([hd[:, 0:1] / ((w - 1.0) / 2.0), hd[:, 1:2] / ((h - 1.0) / 2.0)], dim=1)
        wv = F.grid_sample(prev, grid + fl.permute(0, 2, 3, 1), padding_mode="border", align_corners=False)
        m = torch.sigmoid(hd[:, 2 + c:3 + c])
        warped.copy_(wv)
        pred.copy_(m * wv + (1 - m) * hd[:, 2:2 + c])

    def nan_flag(self, x, flag, bit):
        self.launches += 1
        if bool(torch.isnan(x).any()):
            flag |= bit

    # ------------------------------------------------------------------ layout
    def nchw_to_nhwc(self, x, out):
        self.launches += 1
        out.copy_(x.movedim(-3, -1))
        return out

    def nhwc_to_nchw(self, x, out):
        self.launches += 1
        out.copy_(x.movedim(-1, -3))
        return out

    def launch_count(self):
        return self.launches
