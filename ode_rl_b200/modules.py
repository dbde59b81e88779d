"""Drop-in mirrors of the reference's four hot-path modules -- same constructor / forward signatures, attribute
names and parameter (state_dict) names, so models/ODEConvGRU.py and models/VidODE.py use them unchanged:

  DiffEqSolver    modules/DiffEqSolver.py:12-52
  ODEFunc         modules/DiffEqSolver.py:57-80   (gradient_net stays a real nn.Sequential of nn.Conv2d)
  ODEConvGRUCell  modules/ODEConvGRUCell.py:9-78
  ConvGRUCell     modules/ConvGRUCell.py:11-86
  create_convnet  helpers/utils.py:158-183

Parameters live in ordinary nn.Conv2d / nn.GroupNorm modules (checkpoints and optimizers see no difference);
forward() never calls them -- it hands their tensors to the libodex kernels through ops.py / solver.py.
Tensors cross the boundary in the reference's logical layout ([B,C,H,W], time-first [T,B,C,H,W]); results are
channels-last in memory (a zero-copy view of the kernels' NHWC buffers).
"""
from __future__ import annotations

from typing import List, Optional

import weakref

import torch
import torch.nn as nn

from . import kernels as K
from .kernels import ACT_NONE, CONV_TC
from . import ops as _ops
from .ops import ConvNetOp, GradSink, GruOp, _wgrad_f16, from_nhwc, live, live_input, to_nhwc
from .solver import FIXED, odeint_dopri5, odeint_fixed


def create_convnet(n_inputs, n_outputs, n_layers=1, n_units=128, downsize=False, nonlinear='tanh', final_act=True):
    """helpers/utils.py:158-183: Conv3x3, n_layers x [act, Conv3x3], act, Conv3x3 (, Tanh)."""
    if nonlinear == 'tanh':
        act = nn.Tanh
    elif nonlinear == 'relu':
        act = nn.ReLU
    else:
        raise NotImplementedError('Wrong activation function')
    layers: List[nn.Module] = [nn.Conv2d(n_inputs, n_units, 3, 1, 1, dilation=1)]
    for _ in range(n_layers):
        layers.append(act())
        if downsize is False:
            layers.append(nn.Conv2d(n_units, n_units, 3, 1, 1, dilation=1))
        else:
            layers.append(nn.Conv2d(n_units, n_units, 4, 2, 1, dilation=1))
    layers.append(act())
    layers.append(nn.Conv2d(n_units, n_outputs, 3, 1, 1, dilation=1))
    if final_act is True:
        layers.append(nn.Tanh())
    return nn.Sequential(*layers)


def _empty_result(shape, like: torch.Tensor, params) -> torch.Tensor:
    """Result for an EMPTY batch (a ragged shard may hold no sample): zeros of `shape`, wired into autograd so that
    backward() hands zero gradients to the inputs and parameters, as torch's own conv does for a 0-sample batch.
    No kernel runs (the C ABI rejects n = 0)."""
    out = torch.zeros(shape, dtype=torch.float32, device=like.device)
    for t in [like] + list(params):
        if t is not None and t.requires_grad:
            out = out + t.sum() * 0.0
    return out


class _ConvNetFn(torch.autograd.Function):
    """y = f(x) for a conv stack, logical NCHW in/out."""

    @staticmethod
    def forward(ctx, op: ConvNetOp, x: torch.Tensor, *params):
        K.backend().require(x)
        xn = to_nhwc(x.detach().float())
        save = any(ctx.needs_input_grad)
        if save:
            y, saved = op.fwd(xn, save=True)
            ctx.op, ctx.saved, ctx.params = op, saved, params
        else:
            y = op.fwd(xn)
        return from_nhwc(y)

    @staticmethod
    def backward(ctx, dy):
        d = to_nhwc(dy.float())
        if not d.is_contiguous():
            d = d.contiguous()
        sink = GradSink()
        dx = ctx.op.bwd(ctx.saved, d, sink, need_dx=ctx.needs_input_grad[1])
        return (None, from_nhwc(dx) if dx is not None else None) + sink.result(ctx.params)


# ---------------------------------------------------------------------------------------------------------------------
# ConvTranspose2d(k=4, s=2, p=1) on the tensor-core convolution (the first layer of ConvDecode,
# models/ODEConvGRU.py:122-140, is 60 % of the torch-side time of a training step under exact-fp32 cuDNN).
# out[co, 2Y+a, 2X+b] depends on the 2x2 inputs around (Y, X) selected by the phase (a, b): the four phases together are
# ONE 3x3 "same" convolution Cin -> 4*Cout on the input grid (channel = (a, b, co); 4 of the 9 taps of each phase are
# non-zero) followed by a depth-to-space reshuffle.  kh = 3 - 2r + a (r the 3x3 row tap), same for kw / s / b.
_CT_K = ((3, 1, -1), (-1, 2, 0))            # _CT_K[a][r] = kh (or -1: the tap does not reach this phase)


def _ct_index(device):
    idx = torch.zeros(2, 2, 3, 3, dtype=torch.long)
    msk = torch.zeros(2, 2, 3, 3)
    for a in range(2):
        for b in range(2):
            for r in range(3):
                for q in range(3):
                    kh, kw = _CT_K[a][r], _CT_K[b][q]
                    if kh >= 0 and kw >= 0:
                        idx[a, b, r, q] = kh * 4 + kw
                        msk[a, b, r, q] = 1.0
    return idx.reshape(-1).to(device), msk.reshape(1, 1, -1).to(device)


_CT_CACHE: dict = {}


def _ct_w3(wt: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d weight [Cin, Cout, 4, 4] -> the equivalent 3x3 correlation weight [4*Cout, Cin, 3, 3]."""
    key = str(wt.device)
    if key not in _CT_CACHE:
        _CT_CACHE[key] = _ct_index(wt.device)
    idx, msk = _CT_CACHE[key]
    cin, cout = wt.shape[0], wt.shape[1]
    g = wt.detach().permute(1, 0, 2, 3).reshape(cout, cin, 16)[:, :, idx] * msk          # [Cout, Cin, (a,b,r,s)]
    return g.reshape(cout, cin, 2, 2, 3, 3).permute(2, 3, 0, 1, 4, 5).reshape(4 * cout, cin, 3, 3).contiguous()


def _ct_dwt(dw3: torch.Tensor, cin: int, cout: int) -> torch.Tensor:
    """Adjoint of _ct_w3: every (kh, kw) of the 4x4 kernel is used by exactly one (phase, tap)."""
    g = dw3.reshape(2, 2, cout, cin, 3, 3)
    sel = ((1, 2), (0, 1), (1, 1), (0, 0))           # kh -> (a, r)
    rows = []
    for kh in range(4):
        a, r = sel[kh]
        rows.append(torch.stack([g[a, sel[kw][0], :, :, r, sel[kw][1]] for kw in range(4)], dim=-1))   # [Cout, Cin, 4]
    return torch.stack(rows, dim=-2).permute(1, 0, 2, 3).contiguous()                                  # [Cin, Cout, 4, 4]


# ---- the 3x3 "same" convolutions of ConvEncode / ConvDecode on the 16x16 cell grid with 64-channel sources (config A): the
# image-resident kind::f16 kernel of the ConvGRU (odex_conv16, two-term fp16 split) instead of the 3xTF32 tile kernel, and --
# since that kernel measures its sources' per-image maxima anyway -- the weight gradient without a max pass
# (odex_conv2d_wgrad_f16_scaled).  Other shapes (VidODE's 32x32 ... 128x128) stay on odex_conv2d / odex_conv2d_wgrad_packed.
def _c16_ok(be, x: torch.Tensor, n_in: int, n_out: int) -> bool:
    return (_ops.CONV_SPLIT == "f16" and _ops.CONV_IMPL != "simt" and hasattr(be, "conv16") and x.dim() == 4
            and tuple(x.shape[1:3]) == (16, 16) and x.shape[3] == n_in and n_in in (64, 128) and n_out % 64 == 0
            and be.conv16_shape_ok(x[..., :64], x[..., 64:] if n_in == 128 else None, n_out, 3))


def _conv3(be, x: torch.Tensor, w3: torch.Tensor, dgrad: bool, n_out: int, bias, out: torch.Tensor, want_amax: bool):
    """out = conv3x3(x; w3) (dgrad: with the transposed operator of w3) + bias.  Returns the per-image max |x| [n] when the
    kind::f16 kernel ran and want_amax, else None."""
    n_in = x.shape[3]
    if _c16_ok(be, x, n_in, n_out):
        amax = torch.empty((x.shape[0],), dtype=torch.float32, device=x.device) if want_amax and _ops._wgrad_scaled() else None
        be.conv16(x[..., :64], x[..., 64:] if n_in == 128 else None, be.pack_weights(w3, dgrad=dgrad, tc=2), n_out, 3, bias,
                  ACT_NONE, out, **({"amax_out": amax} if amax is not None else {}))
        return amax
    be.conv2d(x, None, None, be.pack_weights(w3, dgrad=dgrad, tc=True), n_out, 3, bias, ACT_NONE, out, impl=CONV_TC)
    return None


def _conv3_wgrad(be, x: torch.Tensor, dy: torch.Tensor, dwp: torch.Tensor, db, amax_x, amax_dy) -> None:
    both = amax_x is not None and amax_dy is not None
    be.conv2d_wgrad_packed(x, None, dy, 3, dwp, db, f16=_wgrad_f16(), **({"amax": (amax_x, amax_dy)} if both else {}))


class _ConvT4x4s2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        be = K.backend()
        be.require(x)
        xn = to_nhwc(x.detach().float())                      # [n, H, W, Cin], zero-copy for a channels-last view
        if not xn.is_contiguous():
            xn = xn.contiguous()
        n, H, W, cin = xn.shape
        cout = weight.shape[1]
        w3 = _ct_w3(weight.float())
        b3 = bias.detach().float().repeat(4) if bias is not None else None
        P = torch.empty((n, H, W, 4 * cout), dtype=torch.float32, device=xn.device)
        ctx.amax_x = _conv3(be, xn, w3, False, 4 * cout, b3, P, True)
        ctx.save_for_backward(xn, w3)
        ctx.has_bias = bias is not None
        return P                                                # phase-blocked: P[n, Y, X, (a*2+b)*Cout + co]

    @staticmethod
    def backward(ctx, dP):
        be = K.backend()
        xn, w3 = ctx.saved_tensors
        n, H, W, cin = xn.shape
        cout = w3.shape[0] // 4
        dP = dP.float().contiguous()
        dx = dw = db = None
        amax_dy = None
        if ctx.needs_input_grad[0]:
            dxn = torch.empty((n, H, W, cin), dtype=torch.float32, device=xn.device)
            amax_dy = _conv3(be, dP, w3, True, cin, None, dxn, ctx.amax_x is not None)
            dx = from_nhwc(dxn)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dwp = torch.zeros((9, 4 * cout, cin), dtype=torch.float32, device=xn.device)
            db3 = torch.zeros((4 * cout,), dtype=torch.float32, device=xn.device)
            _conv3_wgrad(be, xn, dP, dwp, db3, ctx.amax_x, amax_dy)
            dw3 = torch.zeros((4 * cout, cin, 3, 3), dtype=torch.float32, device=xn.device)
            be.wgrad_unpack(dwp, dw3)
            dw = _ct_dwt(dw3, cin, cout)
            if ctx.has_bias:
                db = db3.view(4, cout).sum(0)
        return dx, dw, db


# Conv2d(k=3, s=2, p=1) the same way (ConvEncode's second layer, models/ODEConvGRU.py:101-119): space-to-depth the input
# (cell (Y, X), channel (a, b, c) = in[c, 2Y+a, 2X+b]); input row 2Y + kh - 1 is cell Y-1 / phase 1 for kh = 0 and cell Y /
# phase kh-1 otherwise, so the strided convolution is a 3x3 "same" convolution 4*Cin -> Cout on the cell grid whose r = 2
# row / s = 2 column taps are zero.  The result is produced directly in NHWC, which is what the encoder recurrence reads.
# Conv2d(k=4, s=2, p=1) (the VidODE encoder's down-sampling layers, models/VidODE.py:17-19) is the same mapping with one more
# tap: input row 2Y + kh - 1 for kh = 3 is cell Y+1 / phase 0.
_CS_K = {3: {(0, 1): 0, (1, 0): 1, (1, 1): 2},                 # (tap r, phase a) -> kh
         4: {(0, 1): 0, (1, 0): 1, (1, 1): 2, (2, 0): 3}}


def _cs_index(device, k):
    tab = _CS_K[k]
    idx = torch.zeros(2, 2, 3, 3, dtype=torch.long)
    msk = torch.zeros(2, 2, 3, 3)
    for a in range(2):
        for b in range(2):
            for r in range(3):
                for q in range(3):
                    if (r, a) in tab and (q, b) in tab:
                        idx[a, b, r, q] = tab[(r, a)] * k + tab[(q, b)]
                        msk[a, b, r, q] = 1.0
    return idx.reshape(-1).to(device), msk.reshape(1, 1, -1).to(device)


_CS_CACHE: dict = {}


def _cs_w3(w: torch.Tensor) -> torch.Tensor:
    """Conv2d weight [Cout, Cin, k, k] (k = 3 or 4, stride 2, padding 1) -> 3x3 stride-1 weight [Cout, 4*Cin, 3, 3] over
    space-to-depth cells."""
    k = w.shape[-1]
    key = (str(w.device), k)
    if key not in _CS_CACHE:
        _CS_CACHE[key] = _cs_index(w.device, k)
    idx, msk = _CS_CACHE[key]
    cout, cin = w.shape[0], w.shape[1]
    g = w.detach().reshape(cout, cin, k * k)[:, :, idx] * msk                   # [Cout, Cin, (a,b,r,s)]
    return g.reshape(cout, cin, 2, 2, 3, 3).permute(0, 2, 3, 1, 4, 5).reshape(cout, 4 * cin, 3, 3).contiguous()


def _cs_dw(dw3: torch.Tensor, cout: int, cin: int, k: int = 3) -> torch.Tensor:
    g = dw3.reshape(cout, 2, 2, cin, 3, 3)
    sel = sorted(_CS_K[k], key=_CS_K[k].get)         # kh -> (r, a)
    rows = []
    for kh in range(k):
        r, a = sel[kh]
        rows.append(torch.stack([g[:, a, sel[kw][1], :, r, sel[kw][0]] for kw in range(k)], dim=-1))   # [Cout, Cin, k]
    return torch.stack(rows, dim=-2).contiguous()                                                       # [Cout, Cin, k, k]


class _Conv3x3s2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        be = K.backend()
        be.require(x)
        n, cin, H2, W2 = x.shape
        H, W = H2 // 2, W2 // 2
        cout = weight.shape[0]
        xs = x.detach().float().reshape(n, cin, H, 2, W, 2).permute(0, 2, 4, 3, 5, 1).reshape(n, H, W, 4 * cin).contiguous()
        w3 = _cs_w3(weight.float())
        out = torch.empty((n, H, W, cout), dtype=torch.float32, device=x.device)
        ctx.amax_x = _conv3(be, xs, w3, False, cout, bias.detach().float() if bias is not None else None, out, True)
        ctx.save_for_backward(xs, w3)
        ctx.has_bias = bias is not None
        ctx.cin, ctx.k = cin, weight.shape[-1]
        return from_nhwc(out)                                   # logical [n, Cout, H, W], channels-last memory

    @staticmethod
    def backward(ctx, dout):
        be = K.backend()
        xs, w3 = ctx.saved_tensors
        n, H, W, c4 = xs.shape
        cin, cout = ctx.cin, w3.shape[0]
        dy = to_nhwc(dout.float())
        if not dy.is_contiguous():
            dy = dy.contiguous()
        dx = dw = db = None
        amax_dy = None
        if ctx.needs_input_grad[0]:
            dxs = torch.empty((n, H, W, c4), dtype=torch.float32, device=xs.device)
            amax_dy = _conv3(be, dy, w3, True, c4, None, dxs, ctx.amax_x is not None)
            dx = dxs.view(n, H, W, 2, 2, cin).permute(0, 5, 1, 3, 2, 4).reshape(n, cin, 2 * H, 2 * W)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dwp = torch.zeros((9, cout, c4), dtype=torch.float32, device=xs.device)
            dbv = torch.zeros((cout,), dtype=torch.float32, device=xs.device)
            _conv3_wgrad(be, xs, dy, dwp, dbv, ctx.amax_x, amax_dy)
            dw3 = torch.zeros((cout, c4, 3, 3), dtype=torch.float32, device=xs.device)
            be.wgrad_unpack(dwp, dw3)
            dw = _cs_dw(dw3, cout, cin, ctx.k)
            if ctx.has_bias:
                db = dbv
        return dx, dw, db


class _ConvCellsFn(torch.autograd.Function):
    """The strided convolution of _Conv3x3s2Fn on an input that already IS in space-to-depth cells [n, H, W, 4*Cin] (what
    odex_encode_stem_fwd writes): no shuffle on the way in, and the data gradient goes back in the same layout."""

    @staticmethod
    def forward(ctx, xs: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        be = K.backend()
        be.require(xs)
        n, H, W, c4 = xs.shape
        cout = weight.shape[0]
        xs = xs.detach()
        w3 = _cs_w3(weight.float())
        out = torch.empty((n, H, W, cout), dtype=torch.float32, device=xs.device)
        ctx.amax_x = _conv3(be, xs, w3, False, cout, bias.detach().float() if bias is not None else None, out, True)
        ctx.save_for_backward(xs, w3)
        ctx.has_bias = bias is not None
        ctx.cin, ctx.k = c4 // 4, weight.shape[-1]
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dout):
        be = K.backend()
        xs, w3 = ctx.saved_tensors
        n, H, W, c4 = xs.shape
        cin, cout = ctx.cin, w3.shape[0]
        dy = to_nhwc(dout.float())
        if not dy.is_contiguous():
            dy = dy.contiguous()
        dxs = dw = db = None
        amax_dy = None
        if ctx.needs_input_grad[0]:
            dxs = torch.empty((n, H, W, c4), dtype=torch.float32, device=xs.device)
            amax_dy = _conv3(be, dy, w3, True, c4, None, dxs, ctx.amax_x is not None)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dwp = torch.zeros((9, cout, c4), dtype=torch.float32, device=xs.device)
            dbv = torch.zeros((cout,), dtype=torch.float32, device=xs.device)
            _conv3_wgrad(be, xs, dy, dwp, dbv, ctx.amax_x, amax_dy)
            dw3 = torch.zeros((cout, c4, 3, 3), dtype=torch.float32, device=xs.device)
            be.wgrad_unpack(dwp, dw3)
            dw = _cs_dw(dw3, cout, cin, ctx.k)
            if ctx.has_bias:
                db = dbv
        return dxs, dw, db


class _EncodeStemFn(torch.autograd.Function):
    """Conv2d(1, 16, 3, 2, 1) + LeakyReLU on [n,1,64,64] frames -> space-to-depth cells [n,16,16,64] (odex_encode_stem_fwd);
    backward gives dw / db only: the frames are data."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], slope: float):
        be = K.backend()
        be.require(x)
        xc = x.detach().float().contiguous()
        w = weight.detach().float().reshape(16, 9).contiguous()
        cells = torch.empty((x.shape[0], 16, 16, 64), dtype=torch.float32, device=x.device)
        be.encode_stem_fwd(xc, w, bias.detach().float() if bias is not None else None, slope, cells)
        ctx.save_for_backward(xc, cells)
        ctx.slope, ctx.has_bias = slope, bias is not None
        return cells

    @staticmethod
    def backward(ctx, dcells):
        if ctx.needs_input_grad[0]:
            raise RuntimeError("encode_stem: the frames do not get a gradient on this path (use the torch layers)")
        xc, cells = ctx.saved_tensors
        d = dcells.float()
        if not d.is_contiguous():
            d = d.contiguous()
        dw = torch.zeros((16, 9), dtype=torch.float32, device=xc.device)
        db = torch.zeros((16,), dtype=torch.float32, device=xc.device) if ctx.has_bias else None
        K.backend().encode_stem_bwd(xc, cells, d, ctx.slope, dw, db)
        return None, dw.view(16, 1, 3, 3), db, None


def encode_stem_supported(x: torch.Tensor, conv: nn.Module, act: nn.Module, nxt: nn.Module) -> bool:
    """True when (conv, act) are ConvEncode's Conv2d(1, 16, 3, 2, 1) + LeakyReLU on 64x64 frames that need no gradient and
    `nxt` is a strided convolution the tensor-core path takes on the resulting 16x16 cells."""
    if not (isinstance(conv, nn.Conv2d) and isinstance(act, nn.LeakyReLU) and isinstance(nxt, nn.Conv2d)):
        return False
    if not (conv.in_channels == 1 and conv.kernel_size == (3, 3) and conv.stride == (2, 2) and conv.padding == (1, 1)
            and conv.dilation == (1, 1) and conv.padding_mode == "zeros" and act.negative_slope > 0):
        return False
    if not (x.is_cuda and x.dtype == torch.float32 and not x.requires_grad
            and K.backend().encode_stem_supported(x, conv.out_channels)):
        return False
    return nxt.in_channels == conv.out_channels and nxt.kernel_size in ((3, 3), (4, 4)) and nxt.stride == (2, 2) \
        and nxt.padding == (1, 1) and nxt.dilation == (1, 1) and nxt.groups == 1 and nxt.padding_mode == "zeros" \
        and K.tc_shape_ok(nxt.out_channels, 4 * nxt.in_channels, 0, 16, 16, 3) \
        and K.tc_shape_ok(4 * nxt.in_channels, nxt.out_channels, 0, 16, 16, 3)


def encode_stem_then_conv(x: torch.Tensor, conv: nn.Conv2d, act: nn.LeakyReLU, nxt: nn.Conv2d) -> torch.Tensor:
    """nxt(act(conv(x))) for ConvEncode's first two layers: the stem kernel writes the activated 16-channel map as
    space-to-depth cells, which the strided tensor-core convolution reads directly (no NCHW intermediate, no shuffle)."""
    cells = _EncodeStemFn.apply(x, conv.weight, conv.bias, float(act.negative_slope))
    return _ConvCellsFn.apply(cells, nxt.weight, nxt.bias)


def conv3x3s2_supported(x: torch.Tensor, m: nn.Module) -> bool:
    """True when `m` is a Conv2d(k=3 or 4, s=2, p=1) that the tensor-core convolution can take for input x."""
    if not (isinstance(m, nn.Conv2d) and m.kernel_size in ((3, 3), (4, 4)) and m.stride == (2, 2) and m.padding == (1, 1)
            and m.dilation == (1, 1) and m.groups == 1 and m.padding_mode == "zeros"):
        return False
    if not x.is_cuda or x.dim() != 4 or x.shape[0] == 0 or x.shape[2] % 2 or x.shape[3] % 2:
        return False
    return (4 * m.in_channels) % 64 == 0 and K.tc_shape_ok(m.out_channels, 4 * m.in_channels, 0, x.shape[2] // 2, x.shape[3] // 2, 3) \
        and K.tc_shape_ok(4 * m.in_channels, m.out_channels, 0, x.shape[2] // 2, x.shape[3] // 2, 3)


def conv3x3s2(x: torch.Tensor, m: nn.Conv2d) -> torch.Tensor:
    """F.conv2d(x, m.weight, m.bias, stride=2, padding=1) for a 3x3 or 4x4 kernel on the 3xTF32 tensor-core path; the result is a
    channels-last view (its memory is the NHWC tensor the encoder recurrence reads without a transpose)."""
    return _Conv3x3s2Fn.apply(x, m.weight, m.bias)


def conv_transpose4x4s2_supported(x: torch.Tensor, m: nn.Module) -> bool:
    """True when `m` is a ConvTranspose2d(k=4, s=2, p=1) that the tensor-core convolution can take for input x."""
    if not (isinstance(m, nn.ConvTranspose2d) and m.kernel_size == (4, 4) and m.stride == (2, 2) and m.padding == (1, 1)
            and m.output_padding == (0, 0) and m.dilation == (1, 1) and m.groups == 1):
        return False
    if not x.is_cuda or x.dim() != 4 or x.shape[0] == 0:
        return False
    return K.tc_shape_ok(4 * m.out_channels, m.in_channels, 0, x.shape[2], x.shape[3], 3)


def conv_transpose4x4s2_phases(x: torch.Tensor, m: nn.ConvTranspose2d) -> torch.Tensor:
    """The transposed convolution in phase-blocked form: P[n, Y, X, (a*2+b)*Cout + co] = out[n, co, 2Y+a, 2X+b]."""
    return _ConvT4x4s2Fn.apply(x, m.weight, m.bias)


def phases_to_nchw(P: torch.Tensor) -> torch.Tensor:
    """depth-to-space of the phase-blocked tensor: [n, H, W, 4*C] -> [n, C, 2H, 2W] (torch ops; autograd handles them)."""
    n, H, W, c4 = P.shape
    c = c4 // 4
    return P.view(n, H, W, 2, 2, c).permute(0, 5, 1, 3, 2, 4).reshape(n, c, 2 * H, 2 * W)


def conv_transpose4x4s2(x: torch.Tensor, m: nn.ConvTranspose2d) -> torch.Tensor:
    """F.conv_transpose2d(x, m.weight, m.bias, stride=2, padding=1) for a 4x4 kernel on the 3xTF32 tensor-core path
    (same parameters, same result to fp32 rounding, hand-written backward)."""
    return phases_to_nchw(conv_transpose4x4s2_phases(x, m))


class _DecodeTailFn(torch.autograd.Function):
    """LeakyReLU + ConvTranspose2d(C, 1, 4, 2, 1) on the phase-blocked tensor (odex_decode_tail_fwd / _bwd)."""

    @staticmethod
    def forward(ctx, P: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], slope: float):
        be = K.backend()
        Pd = P.detach()
        n, H, W, _ = Pd.shape
        w = weight.detach().float().reshape(weight.shape[0], 16).contiguous()
        out = torch.empty((n, 4 * H, 4 * W), dtype=torch.float32, device=Pd.device)
        be.decode_tail_fwd(Pd, w, bias.detach().float() if bias is not None else None, slope, out)
        ctx.save_for_backward(Pd, w)
        ctx.slope, ctx.has_bias, ctx.wshape = slope, bias is not None, tuple(weight.shape)
        return out.view(n, 1, 4 * H, 4 * W)

    @staticmethod
    def backward(ctx, dout):
        be = K.backend()
        Pd, w = ctx.saved_tensors
        dout = dout.float().contiguous()
        dP = torch.empty_like(Pd)
        need_w = ctx.needs_input_grad[1]
        need_b = ctx.has_bias and ctx.needs_input_grad[2]
        dw = torch.zeros_like(w) if need_w else None
        db = torch.zeros((1,), dtype=torch.float32, device=Pd.device) if need_b else None
        be.decode_tail_bwd(Pd, dout, w, ctx.slope, dP, dw, db)
        return dP, (dw.view(ctx.wshape) if need_w else None), db, None


def decode_tail_supported(P: torch.Tensor, act: nn.Module, m: nn.Module) -> bool:
    """True when (act, m) = (LeakyReLU, ConvTranspose2d(C, 1, 4, 2, 1)) can run as the fused decoder tail on P."""
    if not (isinstance(act, nn.LeakyReLU) and isinstance(m, nn.ConvTranspose2d) and m.kernel_size == (4, 4)
            and m.stride == (2, 2) and m.padding == (1, 1) and m.output_padding == (0, 0) and m.dilation == (1, 1)
            and m.groups == 1 and m.out_channels == 1 and 4 * m.in_channels == P.shape[-1]):
        return False
    be = K.backend()
    return hasattr(be, "decode_tail_fwd") and be.decode_tail_supported(P)


def decode_tail(P: torch.Tensor, act: nn.LeakyReLU, m: nn.ConvTranspose2d) -> torch.Tensor:
    """m(act(depth_to_space(P))) in one kernel each way."""
    return _DecodeTailFn.apply(P, m.weight, m.bias, float(act.negative_slope))


# ---------------------------------------------------------------------------------------------------------------------
# A lone "same" Conv2d (3x3 / 5x5 / 1x1, stride 1) of a torch-side caller on the path's tensor-core convolution: the
# per-step decoder of models/VidODE.py:28-44 (conv3x3 256 -> 128 on 64x64 and 128 -> 64 on 128x128 are 78 % of the VidODE
# forward under exact-fp32 cuDNN).  Same parameters (the module's own nn.Conv2d), fp32-level result,
# forward and backward through ConvNetOp like the ODE function's layers.
_LONE_OPS: "weakref.WeakKeyDictionary[nn.Conv2d, ConvNetOp]" = weakref.WeakKeyDictionary()


def conv_same_supported(x: torch.Tensor, m: nn.Module) -> bool:
    if not (isinstance(m, nn.Conv2d) and x.is_cuda and x.dim() == 4 and x.dtype == torch.float32 and x.shape[0] > 0):
        return False
    k = m.kernel_size
    if not (k[0] == k[1] and k[0] in (1, 3, 5) and m.stride == (1, 1) and m.dilation == (1, 1) and m.groups == 1
            and m.padding == (k[0] // 2, k[0] // 2) and m.padding_mode == "zeros"):
        return False
    return K.tc_shape_ok(m.out_channels, m.in_channels, 0, x.shape[2], x.shape[3], k[0])


def conv_same(x: torch.Tensor, m: nn.Conv2d) -> torch.Tensor:
    """m(x) on odex_conv2d (3xTF32 tcgen05); logical NCHW in / out, the result is a channels-last view."""
    op = _LONE_OPS.get(m)
    if op is None:
        op = _LONE_OPS[m] = ConvNetOp(nn.Sequential(m), "conv_same")
    return _ConvNetFn.apply(op, live_input(x), *live(op.params()))


class _ConvSameWFn(torch.autograd.Function):
    """F.conv2d(x, weight, bias, padding=k//2) with the weight / bias given as tensors (so the caller may pad or slice them
    under autograd), on odex_conv2d / odex_conv2d_wgrad_packed; logical NCHW in / out on NHWC memory."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        be = K.backend()
        be.require(x)
        xn = to_nhwc(x.detach().float())
        if not xn.is_contiguous():
            xn = xn.contiguous()
        n, H, W, cin = xn.shape
        cout, R = weight.shape[0], weight.shape[-1]
        w = weight.detach().float().contiguous()
        out = torch.empty((n, H, W, cout), dtype=torch.float32, device=x.device)
        be.conv2d(xn, None, None, be.pack_weights(w, dgrad=False, tc=True), cout, R,
                  bias.detach().float() if bias is not None else None, ACT_NONE, out, impl=CONV_TC)
        ctx.save_for_backward(xn, w)
        ctx.has_bias = bias is not None
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dout):
        be = K.backend()
        xn, w = ctx.saved_tensors
        n, H, W, cin = xn.shape
        cout, R = w.shape[0], w.shape[-1]
        dy = to_nhwc(dout.float())
        if not dy.is_contiguous():
            dy = dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dxn = torch.empty((n, H, W, cin), dtype=torch.float32, device=xn.device)
            be.conv2d(dy, None, None, be.pack_weights(w, dgrad=True, tc=True), cin, R, None, ACT_NONE, dxn, impl=CONV_TC)
            dx = from_nhwc(dxn)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dwp = torch.zeros((R * R, cout, cin), dtype=torch.float32, device=xn.device)
            dbv = torch.zeros((cout,), dtype=torch.float32, device=xn.device)
            be.conv2d_wgrad_packed(xn, None, dy, R, dwp, dbv, f16=_wgrad_f16())
            dw = torch.zeros((cout, cin, R, R), dtype=torch.float32, device=xn.device)
            be.wgrad_unpack(dwp, dw)
            if ctx.has_bias:
                db = dbv
        return dx, dw, db


def conv_same_padded_supported(x: torch.Tensor, m: nn.Module) -> bool:
    """A "same" Conv2d with FEWER than 64 output channels (the VidODE decoder's 64 -> in_channels + 3 head) that fits the
    tensor-core convolution once its output channels are zero-padded to 64."""
    if not (isinstance(m, nn.Conv2d) and x.is_cuda and x.dim() == 4 and x.dtype == torch.float32 and x.shape[0] > 0):
        return False
    k = m.kernel_size
    if not (k[0] == k[1] and k[0] in (1, 3, 5) and m.stride == (1, 1) and m.dilation == (1, 1) and m.groups == 1
            and m.padding == (k[0] // 2, k[0] // 2) and m.padding_mode == "zeros" and m.out_channels < 64):
        return False
    return K.tc_shape_ok(64, m.in_channels, 0, x.shape[2], x.shape[3], k[0]) and \
        K.tc_shape_ok(m.in_channels, 64, 0, x.shape[2], x.shape[3], k[0])


def conv_same_padded(x: torch.Tensor, m: nn.Conv2d) -> torch.Tensor:
    """m(x) for a narrow head: weight / bias zero-padded to 64 output channels (torch ops, so autograd slices the gradients
    back), one tensor-core convolution, the first out_channels channels returned.  cuDNN's exact-fp32 kernels run such a
    6-channel head at ~6 TFLOP/s; the padded convolution costs 10x the FLOPs and still a fifth of the time."""
    pad = 64 - m.out_channels
    w = nn.functional.pad(m.weight, (0, 0, 0, 0, 0, 0, 0, pad))
    b = nn.functional.pad(m.bias, (0, pad)) if m.bias is not None else None
    return _ConvSameWFn.apply(x, w, b)[:, :m.out_channels]


class _Upsample2xFn(torch.autograd.Function):
    """nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False), logical NCHW in / out on NHWC memory."""

    @staticmethod
    def forward(ctx, x: torch.Tensor):
        be = K.backend()
        be.require(x)
        xn = to_nhwc(x.detach().float())
        n, H, W, C_ = xn.shape
        out = torch.empty((n, 2 * H, 2 * W, C_), dtype=torch.float32, device=x.device)
        be.upsample2x_fwd(xn, out)
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dy):
        d = to_nhwc(dy.float())
        if not d.is_contiguous():
            d = d.contiguous()
        n, H2, W2, C_ = d.shape
        dx = torch.empty((n, H2 // 2, W2 // 2, C_), dtype=torch.float32, device=d.device)
        K.backend().upsample2x_bwd(d, dx)
        return from_nhwc(dx)


def upsample2x_supported(x: torch.Tensor, m: nn.Module) -> bool:
    if not (isinstance(m, nn.Upsample) and x.is_cuda and x.dim() == 4 and x.dtype == torch.float32 and x.shape[0] > 0):
        return False
    sf = m.scale_factor
    sf = tuple(sf) if isinstance(sf, (tuple, list)) else (sf, sf)
    return m.mode == "bilinear" and not m.align_corners and sf == (2, 2) and m.size is None and x.shape[1] % 4 == 0


def upsample2x(x: torch.Tensor) -> torch.Tensor:
    """The VidODE decoder's bilinear 2x upsample (models/VidODE.py:34) as one HBM-bound NHWC kernel (odex_upsample2x_fwd;
    adjoint odex_upsample2x_bwd); the result is a channels-last view, what conv_same reads without a repack."""
    return _Upsample2xFn.apply(x)


class _BnReluFn(torch.autograd.Function):
    """relu(BatchNorm2d(x)) with batch statistics (training mode), logical NCHW in / out on NHWC memory (odex_bn_relu_fwd / _bwd).
    Returns (y, mean, biased var); the caller updates the module's running statistics."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, relu: bool):
        be = K.backend()
        be.require(x)
        xn = to_nhwc(x.detach().float())
        if not xn.is_contiguous():
            xn = xn.contiguous()
        Cc = xn.shape[-1]
        y = torch.empty_like(xn)
        stats = torch.empty((3, Cc), dtype=torch.float32, device=xn.device)
        be.bn_relu_fwd(xn, gamma, beta, eps, relu, y, stats)
        ctx.save_for_backward(xn, y, gamma.detach(), stats)
        ctx.relu = relu
        ctx.mark_non_differentiable(stats)
        return from_nhwc(y), stats

    @staticmethod
    def backward(ctx, dy, _dstats):
        be = K.backend()
        xn, y, gamma, stats = ctx.saved_tensors
        d = to_nhwc(dy.float())
        if not d.is_contiguous():
            d = d.contiguous()
        dx = torch.empty_like(xn)
        dgb = torch.empty((2, xn.shape[-1]), dtype=torch.float32, device=xn.device)
        be.bn_relu_bwd(xn, y, d, gamma, stats, ctx.relu, dx, dgb)
        return from_nhwc(dx), dgb[1], dgb[0], None, None


def bn_relu_supported(x: torch.Tensor, bn: nn.Module) -> bool:
    """BatchNorm2d in training mode (batch statistics, what the reference's VidODE runs) with affine parameters on a CUDA
    fp32 tensor whose channel count the kernel takes."""
    return (isinstance(bn, nn.BatchNorm2d) and bn.training and bn.affine and x.is_cuda and x.dim() == 4 and x.dtype == torch.float32
            and x.shape[0] > 0 and x.shape[1] % 4 == 0 and x.shape[1] <= 256 and hasattr(K.backend(), "bn_relu_fwd"))


def bn_relu(x: torch.Tensor, bn: nn.BatchNorm2d, relu: bool = True) -> torch.Tensor:
    """relu(bn(x)) (models/VidODE.py:16-19,36) as one statistics pass + one apply pass; the running statistics are updated as
    nn.BatchNorm2d does (momentum, unbiased variance)."""
    y, stats = _BnReluFn.apply(x, bn.weight, bn.bias, float(bn.eps), relu)
    if bn.track_running_stats and bn.running_mean is not None:
        with torch.no_grad():
            n = x.numel() // x.shape[1]
            bn.num_batches_tracked.add_(1)
            mom = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
            bn.running_mean.mul_(1.0 - mom).add_(stats[0], alpha=mom)
            bn.running_var.mul_(1.0 - mom).add_(stats[2], alpha=mom * n / max(n - 1, 1))
    return y


class _WarpComposeFn(torch.autograd.Function):
    """One step of the recursive flow warp + mask composition (models/VidODE.py:121-131,162-186): forward on
    odex_warp_compose_fwd; the backward re-derives the step with torch's own grid_sample on the saved operands (the warp is
    0.6 % of the forward and off every benchmarked path)."""

    @staticmethod
    def forward(ctx, prev: torch.Tensor, head: torch.Tensor, lin_x: torch.Tensor, lin_y: torch.Tensor):
        be = K.backend()
        be.require(prev)
        p = prev.detach().float().contiguous()
        hn = head.detach().movedim(1, -1)                       # [B, H, W, 3 + C] view of the head's NHWC memory
        if hn.stride(3) != 1 or hn.stride(2) % 4 != 0 or hn.stride(2) < 8 or hn.data_ptr() % 16 != 0:
            hn = torch.nn.functional.pad(hn, (0, 8 - hn.shape[3] % 8 if hn.shape[3] % 8 else 0)).contiguous()
        warped, pred = torch.empty_like(p), torch.empty_like(p)
        be.warp_compose_fwd(p, hn, lin_x, lin_y, warped, pred)
        ctx.save_for_backward(p, head.detach(), lin_x, lin_y)
        return warped, pred

    @staticmethod
    def backward(ctx, dwarped, dpred):
        p, head, lin_x, lin_y = ctx.saved_tensors
        b, c, h, w = p.shape
        with torch.enable_grad():
            p_ = p.requires_grad_(True)
            hd = head.requires_grad_(True)
            grid = torch.stack(torch.broadcast_tensors(lin_x.view(1, 1, w), lin_y.view(1, h, 1)), dim=-1).expand(b, h, w, 2)
            fl = torch.cat([hd[:, 0:1] / ((w - 1.0) / 2.0), hd[:, 1:2] / ((h - 1.0) / 2.0)], dim=1)
            warped = nn.functional.grid_sample(p_, grid + fl.permute(0, 2, 3, 1), padding_mode="border", align_corners=False)
            m = torch.sigmoid(hd[:, 2 + c:3 + c])
            pred = m * warped + (1 - m) * hd[:, 2:2 + c]
            gp, gh = torch.autograd.grad([warped, pred], [p_, hd], [dwarped, dpred], allow_unused=True)
        return gp, gh, None, None


def warp_compose_supported(prev: torch.Tensor, head: torch.Tensor) -> bool:
    return (prev.is_cuda and prev.dim() == 4 and 1 <= prev.shape[1] <= 3 and head.dim() == 4 and head.shape[1] == prev.shape[1] + 3
            and prev.dtype == torch.float32 and head.dtype == torch.float32 and prev.shape[0] > 0 and hasattr(K.backend(), "warp_compose_fwd"))


def warp_compose(prev: torch.Tensor, head: torch.Tensor, lin_x: torch.Tensor, lin_y: torch.Tensor):
    """(warped, pred) for one predicted frame: prev [B,C,H,W], head [B,3+C,H,W] = (flow_x, flow_y, intermediate, mask logit)."""
    return _WarpComposeFn.apply(prev, head, lin_x, lin_y)


class ODEFunc(nn.Module):
    def __init__(self, n_inputs=3, n_outputs=3, n_layers=4, n_units=256, downsize=False, nonlinear='relu',
                 final_act=True, net=None, device=torch.device("cpu")):
        super(ODEFunc, self).__init__()
        self.device = device
        self.n_outputs = n_outputs
        if net is None:
            self.gradient_net = create_convnet(n_inputs, n_outputs, n_layers, n_units, downsize, nonlinear,
                                               final_act=final_act).to(device)
        else:
            self.gradient_net = net
        self._op: Optional[ConvNetOp] = None

    def op(self) -> ConvNetOp:
        """Kernel-side view of gradient_net (built lazily; raises NotImplementedError for unsupported stacks)."""
        if self._op is None or self._op.layers[0][0] is not self.gradient_net[0]:
            self._op = ConvNetOp(self.gradient_net, "ODEFunc.gradient_net")
        return self._op

    def forward(self, t_local, y, backwards=False):
        """dy/dt at y; autonomous: t_local is ignored (modules/DiffEqSolver.py:71-80)."""
        op = self.op()
        if y.shape[0] == 0:
            grad = _empty_result((0, op.c_out) + tuple(y.shape[2:]), y, op.params())
        else:
            grad = _ConvNetFn.apply(op, live_input(y), *live(op.params()))
        if backwards:
            grad = -grad
        return grad


class DiffEqSolver(nn.Module):
    def __init__(self, ode_func, method, odeint_rtol=1e-4, odeint_atol=1e-5, device=torch.device("cpu"), memory=False):
        super(DiffEqSolver, self).__init__()
        self.ode_func = ode_func
        self.ode_method = method
        self.device = device
        self.memory = memory
        self.odeint_rtol = odeint_rtol
        self.odeint_atol = odeint_atol
        # dopri5 plumbing (not part of the reference surface)
        self.reduce_fn = None            # set by parallel.shard_solver(): all-reduce of the error-norm partial sums
        self.max_attempts = 4096
        self.save_budget_bytes = 8 << 30
        self.last_stats = None
        self._chunk_hint = 8

    def _odeint(self, y0, t):
        method = self.ode_method
        if t.dim() != 1:
            raise AssertionError("t must be one dimensional")
        if not isinstance(self.ode_func, ODEFunc):
            raise NotImplementedError("DiffEqSolver: the CUDA path integrates ode_rl_b200.ODEFunc conv stacks only")
        op = self.ode_func.op()
        if y0.shape[0] == 0:
            if method not in FIXED and method != 'dopri5':
                raise ValueError('Invalid method "{}" (euler, rk4 and dopri5 are supported)'.format(method))
            if method == 'dopri5' and self.reduce_fn is not None:
                raise ValueError("DiffEqSolver(dopri5, sharded): an empty shard cannot take part in the batch-shared step "
                                 "control; give every rank at least one sample")
            return _empty_result((t.numel(), 0) + tuple(y0.shape[1:]), y0, op.params())
        if method in FIXED:
            return odeint_fixed(op, y0, t, method)          # (validates t when that needs no device sync)
        if method == 'dopri5':
            return odeint_dopri5(op, self, y0, t, self.odeint_rtol, self.odeint_atol)
        raise ValueError('Invalid method "{}" (euler, rk4 and dopri5 are supported)'.format(method))

    def forward(self, first_point, time_steps_to_predict, backwards=False):
        """Decode the trajectory through the ODE solver; returns time-first [T,B,C,H,W] (modules/DiffEqSolver.py:24-52)."""
        if self.memory is True:
            # modules/DiffEqSolver.py:30-42: odeint over a single time point returns y0, so h_next = 2*h_prev.
            y_is = [first_point]
            b, c, h, w = first_point.size()
            for i in range(len(time_steps_to_predict)):
                h_prev = y_is[-1]
                pred_m = self._odeint(h_prev, time_steps_to_predict[i:i + 1]).reshape(b, c, h, w)
                y_is.append(h_prev + pred_m)
            return torch.stack(y_is[1:]).permute(1, 0, 2, 3, 4)
        if first_point.dim() != 4:
            # the reference returns None for vector states (modules/DiffEqSolver.py:48-51); nothing to integrate here
            return None
        return self._odeint(first_point, time_steps_to_predict)


class _GruStepFn(torch.autograd.Function):
    """h' = ConvGRU(x, h), logical NCHW in/out (one step of modules/ConvGRUCell.py:66-84)."""

    @staticmethod
    def forward(ctx, op: GruOp, x: torch.Tensor, h: torch.Tensor, *params):
        K.backend().require(x)
        xn, hn = to_nhwc(x.detach().float()), to_nhwc(h.detach().float())
        save = any(ctx.needs_input_grad)
        if save:
            hnew, saved = op.fwd(xn, hn, save=True)
            ctx.op, ctx.saved, ctx.params = op, saved, params
        else:
            hnew = op.fwd(xn, hn)
        return from_nhwc(hnew)

    @staticmethod
    def backward(ctx, dh):
        d = to_nhwc(dh.float())
        if not d.is_contiguous():
            d = d.contiguous()
        op = ctx.op
        sink = GradSink()
        D = op.bwd(ctx.saved, d, sink)
        dx = from_nhwc(D[..., :op.Cx]) if ctx.needs_input_grad[1] else None
        dhp = from_nhwc(D[..., op.Cx:]) if ctx.needs_input_grad[2] else None
        return (None, dx, dhp) + sink.result(ctx.params)


class ConvGRUCell(nn.Module):
    def __init__(self, input_size, input_dim, hidden_dim, kernel_size, bias=True, dtype=None, padding=None):
        """ConvGRU cell with GroupNorm gates -- modules/ConvGRUCell.py:12-50 (same submodule / parameter names)."""
        super(ConvGRUCell, self).__init__()
        self.height, self.width = input_size
        self.input_channels = input_dim
        self.hidden_dim = hidden_dim
        self.device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        if padding is None:
            self.padding = (kernel_size - 1) // 2
        else:
            self.padding = padding
        if dtype is None:
            dtype = torch.cuda.FloatTensor if self.device == 'cuda' else torch.FloatTensor
        self.bias = bias
        self.dtype = dtype
        self.conv_gates = nn.Sequential(
            nn.Conv2d(input_dim + hidden_dim, 2 * self.hidden_dim, kernel_size, 1, self.padding),
            nn.GroupNorm(2 * self.hidden_dim // 32, 2 * self.hidden_dim))
        self.conv_can = nn.Sequential(
            nn.Conv2d(input_dim + hidden_dim, self.hidden_dim, kernel_size, 1, self.padding),
            nn.GroupNorm(self.hidden_dim // 32, self.hidden_dim))
        self._op: Optional[GruOp] = None

    def op(self) -> GruOp:
        if self._op is None or self._op.cg is not self.conv_gates[0]:
            self._op = GruOp(self.conv_gates, self.conv_can, self.hidden_dim, self.input_channels)
        return self._op

    def init_hidden(self, batch_size):
        return (torch.zeros(batch_size, self.hidden_dim, self.height, self.width)).type(self.dtype)

    def forward(self, input_tensor=None, h_cur=None, seq_len=10, mask=None, dim=0):
        """Returns (stack(h_1..h_seq_len, dim), h_seq_len); `mask` is accepted and ignored like the reference."""
        op = self.op()
        dev = self.conv_gates[0].weight.device
        if h_cur is None:
            h_cur = torch.zeros(input_tensor.size(1), self.hidden_dim, self.height, self.width, device=dev)
        output_inner = []
        h_next = h_cur
        for index in range(seq_len):
            if input_tensor is None:
                x = torch.zeros(h_cur.size(0), self.input_channels, self.height, self.width, device=dev)
            else:
                x = input_tensor[index, ...]
            if h_cur.size(0) == 0:
                h_next = _empty_result(tuple(h_cur.shape), h_cur, [x] + op.params())
            else:
                h_next = _GruStepFn.apply(op, live_input(x), live_input(h_cur), *live(op.params()))
            output_inner.append(h_next)
            h_cur = h_next
        return torch.stack(output_inner, dim=dim), h_next


class _EncoderFn(torch.autograd.Function):
    """The ODE-ConvGRU recurrence of modules/ODEConvGRUCell.py:53-74 as one autograd node: for every observed
    frame, walked in `order`, h~ = h + f(h)*dt_i (explicit Euler, fused into the last conv's epilogue) and
    h = ConvGRU(x_i, h~).  Returns all hidden states, iteration-major [T,B,C,H,W], and the final one once more as an output of
    its own [B,C,H,W]: the model reads only that (modules/ODEConvGRUCell.py:32-37), and as a separate output its gradient arrives
    alone -- no [T,B,...] gradient that is zero but for its last slice has to be built, laid out and added step by step."""

    @staticmethod
    def forward(ctx, f_op: ConvNetOp, g_op: GruOp, n_f: int, order, inputs: torch.Tensor, deltas: torch.Tensor, *params):
        ctx.set_materialize_grads(False)
        be = K.backend()
        be.require(inputs)
        x = to_nhwc(inputs.detach().float())                 # [T,B,H,W,C]
        if not x.is_contiguous():
            x = x.contiguous()
        T, B, H, W, Cx = x.shape
        C = g_op.C
        save = any(ctx.needs_input_grad)
        hs = torch.empty((T, B, H, W, C), dtype=torch.float32, device=x.device)
        h = torch.zeros((B, H, W, C), dtype=torch.float32, device=x.device)
        saved = []
        # saved activations of f over the T iterations in [T, B, ...] arenas: the deferred weight gradients of a layer then
        # run as one launch over all iterations (GradSink.add_wgrad)
        arenas = None
        Lf = len(f_op.layers)
        if save and Lf > 1 and all(c.out_channels == C for c, _ in f_op.layers[:-1]):
            arenas = [torch.empty((T, B, H, W, C), dtype=torch.float32, device=x.device) for _ in range(Lf - 1)]
        # per-image operand maxima of f's layers over the T iterations ([L, 2, T, B]): measured by the chain launches, read by the
        # merged weight gradients (ConvNetOp.new_maxima)
        mx_arena = None
        if save and T > 0 and B > 0 and f_op.maxima_ok(h):
            mx_arena = torch.empty((Lf, 2, T, B), dtype=torch.float32, device=x.device)
        # the ConvGRU's weight-gradient operands over the T steps in arenas indexed by FRAME (like x): h~, r*h, and their maxima
        gru_ar = None
        if save and T > 1 and B > 0 and g_op._fused16(be, x[0], h):
            gru_ar = {"ht": torch.empty((T, B, H, W, C), dtype=torch.float32, device=x.device),
                      "rh": torch.empty((T, B, H, W, C), dtype=torch.float32, device=x.device),
                      "mx": torch.empty((2, 2, T, B), dtype=torch.float32, device=x.device)}
        for n, i in enumerate(order):
            ht = gru_ar["ht"][i] if gru_ar is not None else torch.empty_like(h)
            r = f_op.fwd(h, save=save, out=ht, res=h, alpha=1.0, alpha_dev=deltas[n:n + 1],
                         bufs=[a_[n] for a_ in arenas] if arenas is not None else None,
                         maxima=mx_arena[:, :, n] if mx_arena is not None else None)
            g = g_op.fwd(x[i], ht, save=save, out=hs[n],
                         bufs={"rh": gru_ar["rh"][i], "mx": gru_ar["mx"][:, :, i]} if gru_ar is not None else None)
            if save:
                saved.append((r[1], g[1]))
            h = hs[n]
        if save:
            ctx.f_op, ctx.g_op, ctx.n_f, ctx.order, ctx.deltas, ctx.saved, ctx.params = f_op, g_op, n_f, order, deltas, saved, params
            ctx.x_shape = (T, B, H, W, Cx)
            ctx.gru_arenas = gru_ar is not None
        return from_nhwc(hs), from_nhwc(hs[T - 1].clone())

    @staticmethod
    def backward(ctx, dhs, dlast):
        be = K.backend()
        f_op, g_op, order, deltas, saved = ctx.f_op, ctx.g_op, ctx.order, ctx.deltas, ctx.saved
        if saved is None:
            raise RuntimeError("ODEConvGRUCell: this recurrence was already back-propagated through once; its saved "
                               "activations are freed during that pass (no retain_graph) -- run the forward again")

        def dense(g):
            if g is None:
                return None
            g = to_nhwc(g.float())
            return g if g.is_contiguous() else g.contiguous()

        d, dl = dense(dhs), dense(dlast)                      # either may be None: that output was not used
        ref = d if d is not None else dl
        T, B, H, W, Cx = ctx.x_shape
        C = g_op.C
        sink = GradSink()
        sink.deferred = True                                  # weight gradients of f: batched over the iterations at the end
        Lf = len(f_op.layers)
        dev = ref.device
        dk_arena = torch.empty((T, B, H, W, C), dtype=torch.float32, device=dev)
        g_arenas = None
        if Lf > 1 and all(c.in_channels == C for c, _ in f_op.layers[1:]):
            g_arenas = [torch.empty((T, B, H, W, C), dtype=torch.float32, device=dev) for _ in range(Lf - 1)]
        split = bool(ctx.gru_arenas) and Cx % 64 == 0        # odex_conv16 writes dx_i and dh~ into two dense tensors
        if split:
            DX = torch.empty((T, B, H, W, Cx), dtype=torch.float32, device=dev)
            DH = torch.empty((T, B, H, W, C), dtype=torch.float32, device=dev)
        else:
            DX = torch.empty((T, B, H, W, Cx + C), dtype=torch.float32, device=dev)   # [dx_i | dh~] per frame
        gru_bw = None
        if ctx.gru_arenas:                                    # dy of the ConvGRU's weight gradients, indexed by frame like x
            gru_bw = {"dcpre": torch.empty((T, B, H, W, C), dtype=torch.float32, device=dev),
                      "dg": torch.empty((T, B, H, W, 2 * C), dtype=torch.float32, device=dev)}
        a = None                                              # adjoint of h after iteration n
        for n in range(T - 1, -1, -1):
            i = order[n]
            parts = [g_ for g_ in (a, d[n] if d is not None else None, dl if n == T - 1 else None) if g_ is not None]
            if not parts:
                dh = torch.zeros((B, H, W, C), dtype=torch.float32, device=dev)
            elif len(parts) == 1:
                dh = parts[0]
            else:
                dh = torch.empty((B, H, W, C), dtype=torch.float32, device=dev)
                be.lincomb(dh, parts, [1.0] * len(parts))
            sv_f, sv_g = saved[n]
            gb_ = {"dcpre": gru_bw["dcpre"][i], "dg": gru_bw["dg"][i]} if gru_bw is not None else None
            if split:
                _, dht = g_op.bwd(sv_g, dh, sink, bufs=gb_, split_out=(DX[i], DH[i]))      # adjoint of h~, dense
            else:
                D = g_op.bwd(sv_g, dh, sink, out=DX[i], bufs=gb_)
                dht = D[..., Cx:]                             # adjoint of h~ (row stride Cx+C)
            if n == 0:
                # h_0 = 0 is a constant: only the parameter gradients of f(0) remain
                dk = _scaled_copy(be, dht, deltas[n:n + 1], out=dk_arena[n])
                f_op.bwd(sv_f, dk, sink, need_dx=False, gbufs=[g_[n] for g_ in g_arenas] if g_arenas is not None else None)
                a = None
            else:
                dk = _scaled_copy(be, dht, deltas[n:n + 1], out=dk_arena[n])
                a = f_op.bwd(sv_f, dk, sink, res=dht, gbufs=[g_[n] for g_ in g_arenas] if g_arenas is not None else None)
            saved[n] = None
        ctx.saved = None
        dx = from_nhwc(DX if split else DX[..., :Cx]) if ctx.needs_input_grad[4] else None
        return (None, None, None, None, dx, None) + sink.result(ctx.params)


def _scaled_copy(be, view: torch.Tensor, dt_dev: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dense copy of an NHWC channel-slice view scaled by *dt_dev: the adjoint of `f(h) * dt` w.r.t. f(h)."""
    dense = view.contiguous() if not view.is_contiguous() else view
    if out is None:
        out = torch.empty_like(dense)
    be.lincomb(out, [dense], [1.0], dtmask=1, dt_dev=dt_dev)
    return out


class ODEConvGRUCell(nn.Module):
    def __init__(self, ode_func, opt, resolution, ch, out_ch=None, device=None, kernel_size=5):
        super(ODEConvGRUCell, self).__init__()
        self.ode_func = ode_func
        self.device = device
        self.z0_diffeq_solver = None
        if out_ch is None:
            out_ch = ch
        self.cgru_cell = ConvGRUCell(input_size=resolution, input_dim=ch, hidden_dim=ch,
                                     kernel_size=kernel_size, bias=True).to(device)
        # last conv layer for generating mu, sigma
        self.z0_dim = out_ch
        z = ch
        self.transform_z0 = nn.Sequential(
            nn.Conv2d(z, z, 1, 1, 0),
            nn.ReLU(),
            nn.Conv2d(z, out_ch * 2, 1, 1, 0)).to(device)
        self._z0_op: Optional[ConvNetOp] = None
        self._delta_idx: dict = {}
        self._flag: Optional[torch.Tensor] = None     # sticky device error word (bit0: NaN in the recurrence)
        self.check_nan = True

    def _transform_op(self) -> ConvNetOp:
        if self._z0_op is None or self._z0_op.layers[0][0] is not self.transform_z0[0]:
            self._z0_op = ConvNetOp(self.transform_z0, "ODEConvGRUCell.transform_z0")
        return self._z0_op

    def forward(self, inputs, timesteps, mask=None):    # inputs has time dim first
        last_yi, _ = self.run_ode_conv_gru(inputs, timesteps, mask=mask)
        op = self._transform_op()
        if last_yi.shape[0] == 0:
            trans_last_yi = _empty_result((0, op.c_out) + tuple(last_yi.shape[2:]), last_yi, op.params())
        else:
            trans_last_yi = _ConvNetFn.apply(op, live_input(last_yi), *live(op.params()))     # (b, self.z0_dim*2, h, w)
        mean_z0, std_z0 = torch.split(trans_last_yi, self.z0_dim, dim=1)
        std_z0 = std_z0.abs()
        return mean_z0, std_z0

    def _deltas(self, timesteps, device, order):
        """Step sizes exactly as the reference forms them (fp64): prev_t, t_i = t[-1]+0.01, t[-1]; then prev_t, t_i =
        t[i], t[i-1] after visiting frame i (modules/ODEConvGRUCell.py:48,73; index -1 wraps like the reference's).
        Recomputed from the VALUES of `timesteps` on every call, on the device (three tiny torch ops, capturable in a CUDA
        graph); only the gather index -- a function of (order, device) alone -- is cached, because building it is a
        pageable host->device copy, which must not happen during graph capture."""
        t = len(order)
        ts = timesteps.detach().to(device=device, dtype=torch.float64)
        key = (str(device), order)
        idx = self._delta_idx.get(key)
        if idx is None:
            idx = self._delta_idx[key] = (torch.tensor(list(order[:-1]), dtype=torch.long, device=device),
                                          torch.tensor([(i - 1) % t for i in order[:-1]], dtype=torch.long, device=device))
        first = (ts[-1] - (ts[-1] + 0.01)).reshape(1)
        if t == 1:
            return first
        return torch.cat((first, ts[idx[1]] - ts[idx[0]]))

    def check_errors(self) -> None:
        """Raise what the reference's per-frame asserts would have raised (modules/ODEConvGRUCell.py:56,59: NaN in the Euler
        increment / the updated state) from the sticky device word: ONE device->host read, after the forward.  forward()
        calls it itself unless a CUDA graph is being captured or `self.check_nan` is False; a caller that replays a captured
        step calls it whenever it wants the check (e.g. together with reading the loss)."""
        if self._flag is not None and int(self._flag.item()) != 0:
            self._flag.zero_()
            raise AssertionError("NaN in the ODE-ConvGRU encoder state (modules/ODEConvGRUCell.py:56,59)")

    def run_ode_conv_gru(self, inputs, timesteps, run_backwards=True, mask=None):
        t, b, c, h, w = inputs.size()
        assert (t == len(timesteps)), "Sequence length should be same as time_steps"
        if not isinstance(self.ode_func, ODEFunc):
            raise NotImplementedError("ODEConvGRUCell: the CUDA path needs an ode_rl_b200.ODEFunc")
        order = list(range(t))
        if run_backwards:
            order = order[::-1]
        deltas = self._deltas(timesteps, inputs.device, tuple(order))
        f_op, g_op = self.ode_func.op(), self.cgru_cell.op()
        params = f_op.params() + g_op.params()
        h_last = None
        if b == 0:
            hs = _empty_result((t, 0, g_op.C, h, w), inputs, params)
        else:
            hs, h_last = _EncoderFn.apply(f_op, g_op, len(f_op.params()), order, live_input(inputs), deltas, *live(params))   # [T,B,C,H,W], [B,C,H,W]
            if self.check_nan:
                if self._flag is None or self._flag.device != hs.device:
                    self._flag = torch.zeros(1, dtype=torch.int32, device=hs.device)
                last = hs.detach()[t - 1].movedim(-3, -1)            # the kernels' NHWC buffer of the final hidden state
                K.backend().nan_flag(last if last.is_contiguous() else last.contiguous(), self._flag, 1)
                if not (hs.is_cuda and torch.cuda.is_current_stream_capturing()):
                    self.check_errors()
        yi = h_last if h_last is not None else hs[t - 1]
        latent_ys = hs.permute(1, 0, 2, 3, 4)   # (b, t, c, h, w)
        return yi, latent_ys
