"""Tensor-level wrappers over the C-ABI of libodex.so -- the ONLY module that touches ctypes pointers.

Everything above this file (ops.py, solver.py, modules.py) talks to a ``Backend`` object in terms of torch
tensors that merely own device memory; every method below enqueues hand-written sm_100a kernels on the
current CUDA stream and returns immediately (no host sync).  There is no CPU implementation here: calling
any method with the library missing or with non-CUDA tensors raises.

Layout conventions (include/odex.h):
  * activations are fp32 NHWC tensors ``[B, H, W, C']`` whose last dim is dense; a channel slice of a wider
    buffer is passed as a view (its pixel-row stride becomes the kernel's ``ld``);
  * RK planes are dense fp32 tensors of any shape, addressed flat;
  * device scalars (dt, alpha, skip flag) are 1-element views into the solver's state buffer.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import ACT_NONE, ACT_RELU, ACT_TANH, CONV_AUTO, CONV_SIMT, CONV_TC  # noqa: F401 (re-export)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _rows(t: torch.Tensor):
    """(pointer, pixel-row stride) of an NHWC view [B,H,W,C'] that a kernel can address as p*ld + c."""
    assert t.dim() == 4 and t.dtype == torch.float32, "expected an fp32 NHWC tensor [B,H,W,C]"
    B, H, W, Cc = t.shape
    ld = t.stride(2)
    assert t.stride(3) == 1 and (W == 1 or True), "channel dim must be dense"
    assert (H == 1 or t.stride(1) == W * ld) and (B == 1 or t.stride(0) == H * W * ld), \
        "NHWC view must have uniform pixel stride (got strides %s for shape %s)" % (t.stride(), tuple(t.shape))
    return t.data_ptr(), ld


def _flat(t: torch.Tensor) -> int:
    assert t.dtype == torch.float32 and t.is_contiguous(), "RK planes must be dense fp32"
    return t.data_ptr()


class PackedConv:
    """Kernel-ready copies of one Conv2d weight: forward and data-gradient operators."""

    __slots__ = ("N", "C", "R", "fwd_simt", "fwd_tc", "dgrad_simt", "dgrad_tc", "fwd_f16", "dgrad_f16")

    def __init__(self, N, Cin, R):
        self.N, self.C, self.R = N, Cin, R
        self.fwd_simt = self.fwd_tc = self.dgrad_simt = self.dgrad_tc = None
        self.fwd_f16 = self.dgrad_f16 = None         # two-term fp16 split (odex_pack_weights tc = 2)


def tc_shape_ok(N: int, c0: int, c1: int, H: int, W: int, R: int) -> bool:
    """Mirror of conv_tc_supported() in csrc/conv_tc.cu."""
    if (N != 64 and (N % 128 != 0 or N > 1024)) or R not in (1, 3, 5):     # > 128: N-blocks of 128, one launch each
        return False
    if c0 % 64 != 0 or (c1 and c1 != c0):
        return False
    if W < 8 or W > 128 or (W & (W - 1)) != 0:
        return False
    return H % (128 // W) == 0


class Backend:
    """The CUDA backend (libodex.so).  tests/ substitutes a torch-CPU emulation of the same methods to check
    the host-side solver / backward logic without a GPU; the product never does."""

    name = "libodex"

    def __init__(self):
        self.lib = _lib.lib()

    # ------------------------------------------------------------------ memory helpers
    @staticmethod
    def require(t: torch.Tensor):
        if not t.is_cuda:
            raise _lib.OdexError("ode_rl_b200 runs on CUDA tensors only (got %s); there is no CPU fallback" % t.device)

    # ------------------------------------------------------------------ weights
    def pack_weights(self, w: torch.Tensor, dgrad: bool, tc) -> torch.Tensor:
        """odex_pack_weights: OIHW fp32 -> SIMT [tap][c][n] (tc = 0 / False), TC 3xTF32 hi/lo k-blocks (1 / True) or the
        two-term fp16 split with its scale tail (2)."""
        self.require(w)
        w = w.detach()
        assert w.dtype == torch.float32 and w.is_contiguous() and w.dim() == 4 and w.shape[2] == w.shape[3]
        N, Cc, R, _ = w.shape
        n_out, c_in = (Cc, N) if dgrad else (N, Cc)
        mode = int(tc)
        nfl = self.lib.odex_packed_weight_floats(n_out, c_in, R, mode)
        out = torch.empty(nfl, dtype=torch.float32, device=w.device)
        _lib.check(self.lib.odex_pack_weights(w.data_ptr(), N, Cc, R, 1 if dgrad else 0, mode,
                                              out.data_ptr(), _stream()), "odex_pack_weights")
        return out

    def pack_weights_f16_pair(self, w: torch.Tensor):
        """odex_pack_weights_f16_pair: (pack_weights(w, False, 2), pack_weights(w, True, 2)) in one pair of launches."""
        self.require(w)
        w = w.detach()
        assert w.dtype == torch.float32 and w.is_contiguous() and w.dim() == 4 and w.shape[2] == w.shape[3]
        N, Cc, R, _ = w.shape
        nfl = self.lib.odex_packed_weight_floats(N, Cc, R, 2)
        fwd = torch.empty(nfl, dtype=torch.float32, device=w.device)
        dgr = torch.empty(nfl, dtype=torch.float32, device=w.device)
        _lib.check(self.lib.odex_pack_weights_f16_pair(w.data_ptr(), N, Cc, R, fwd.data_ptr(), dgr.data_ptr(), _stream()),
                   "odex_pack_weights_f16_pair")
        return fwd, dgr

    # ------------------------------------------------------------------ convolution
    def conv2d(self, x0: torch.Tensor, x1: Optional[torch.Tensor], w_simt: Optional[torch.Tensor],
               w_tc: Optional[torch.Tensor], N: int, R: int, bias: Optional[torch.Tensor], act: int,
               out: torch.Tensor, dact: Optional[torch.Tensor] = None, dact_kind: int = 0,
               res: Optional[torch.Tensor] = None, alpha: float = 1.0, alpha_dev: Optional[torch.Tensor] = None,
               skip: Optional[torch.Tensor] = None, impl: int = CONV_AUTO, repeat: int = 1) -> torch.Tensor:
        """odex_conv2d: out = res + alpha * (act(conv(cat(x0,x1)) + bias) * act'(dact))."""
        self.require(x0)
        d = _lib.ConvDesc()
        B, H, W, c0 = x0.shape
        d.x0, d.ld0 = _rows(x0)
        d.c0 = c0
        if x1 is not None:
            assert x1.shape[:3] == x0.shape[:3]
            d.x1, d.ld1 = _rows(x1)
            d.c1 = x1.shape[3]
        d.B, d.H, d.W, d.R, d.N = B, H, W, R, N
        d.w_simt, d.w_tc = _ptr(w_simt), _ptr(w_tc)
        d.bias = _ptr(bias.detach()) if bias is not None else None
        d.act = act
        if dact is not None:
            d.dact, d.ldd = _rows(dact)
            d.dact_kind = dact_kind
        if res is not None:
            d.res, d.ldr = _rows(res)
        d.alpha = float(alpha)
        d.alpha_dev = _ptr(alpha_dev)
        assert tuple(out.shape) == (B, H, W, N), "conv2d: out shape %s != %s" % (tuple(out.shape), (B, H, W, N))
        d.out, d.ldo = _rows(out)
        d.skip = _ptr(skip)
        d.impl = impl
        if repeat > 1:
            _lib.check(self.lib.odex_conv2d_repeat(C.byref(d), int(repeat), _stream()), "odex_conv2d_repeat")
        else:
            _lib.check(self.lib.odex_conv2d(C.byref(d), _stream()), "odex_conv2d")
        return out

    def _convnet_desc(self, x, w_tc, bias, act, aux, aux_kind, out, alpha, alpha_dev, skip, split=0, amax_out=None):
        d = _lib.ConvNetDesc()
        d.split = int(split)
        B, H, W, Cc = x.shape
        d.x, d.ldx = _rows(x)
        d.B, d.H, d.W, d.C, d.L = B, H, W, Cc, len(w_tc)
        for l in range(len(w_tc)):
            d.w_tc[l] = _ptr(w_tc[l])
            d.bias[l] = _ptr(bias[l].detach()) if bias[l] is not None else None
            d.act[l] = int(act[l])
            d.aux_kind[l] = int(aux_kind[l]) if aux[l] is not None else 0
            if aux[l] is not None:
                assert tuple(aux[l].shape) == (B, H, W, Cc)
                d.aux[l], d.ld_aux[l] = _rows(aux[l])
            if out[l] is not None:
                assert tuple(out[l].shape) == (B, H, W, Cc)
                d.out[l], d.ld_out[l] = _rows(out[l])
        if amax_out is not None:
            assert split == 1, "amax_out needs the kind::f16 chain"
            for l, t in enumerate(amax_out):
                if t is not None:
                    assert t.dtype == torch.float32 and t.numel() == B and t.is_contiguous() and t.device == x.device
                    d.amax_out[l] = t.data_ptr()
        d.alpha = float(alpha)
        d.alpha_dev = _ptr(alpha_dev)
        d.skip = _ptr(skip)
        return d

    def convnet_supported(self, x: torch.Tensor, n_layers: int) -> bool:
        """Shape test of odex_convnet (16x16 images, 64 -> 64 3x3 layers, at most NET_MAX_LAYERS of them)."""
        B, H, W, Cc = x.shape
        return (H == 16 and W == 16 and Cc == 64 and 1 <= n_layers <= _lib.NET_MAX_LAYERS and B > 0
                and x.stride(3) == 1 and x.stride(2) % 4 == 0 and os.environ.get("ODEX_NO_FUSED_NET") is None)

    def convnet(self, x: torch.Tensor, w_tc: Sequence[torch.Tensor], bias: Sequence[Optional[torch.Tensor]],
                act: Sequence[int], aux: Sequence[Optional[torch.Tensor]], aux_kind: Sequence[int],
                out: Sequence[Optional[torch.Tensor]], alpha: float = 1.0, alpha_dev: Optional[torch.Tensor] = None,
                skip: Optional[torch.Tensor] = None, split: int = 0,
                amax_out: Optional[Sequence[Optional[torch.Tensor]]] = None) -> None:
        """odex_convnet: the whole conv chain in one launch; layer l: v = act(conv(a_l) + bias), v *= act'(aux) (kind 1/2)
        or v = aux + alpha * v (kind 3), stored to out[l] when given (out[-1] required).  split: 0 = w_tc packed with tc = 1
        (3xTF32), 1 = packed with tc = 2 (two-term fp16 split on kind::f16).  amax_out[l] (split = 1): fp32 tensor [B] that
        receives max |input of layer l| per image."""
        self.require(x)
        d = self._convnet_desc(x, w_tc, bias, act, aux, aux_kind, out, alpha, alpha_dev, skip, split, amax_out)
        _lib.check(self.lib.odex_convnet(C.byref(d), _stream()), "odex_convnet")

    # ------------------------------------------------------------------ image-resident conv (+ fused ConvGRU gates), kind::f16
    @staticmethod
    def conv16_shape_ok(x0: torch.Tensor, x1: Optional[torch.Tensor], N: int, R: int) -> bool:
        """Shape test of odex_conv16: 16x16 images, one or two 64-channel sources, N % 64 == 0, R in {3, 5}."""
        if x0.dim() != 4 or tuple(x0.shape[1:]) != (16, 16, 64) or x0.shape[0] == 0 or R not in (3, 5) or N % 64 != 0:
            return False
        if x1 is not None and tuple(x1.shape) != tuple(x0.shape):
            return False
        return os.environ.get("ODEX_NO_CONV16") is None

    def _conv16_desc(self, x0, x1, w_f16, N, R, bias, skip):
        d = _lib.Conv16Desc()
        d.x0, d.ld0 = _rows(x0)
        if x1 is not None:
            d.x1, d.ld1 = _rows(x1)
        d.B, d.R, d.N = x0.shape[0], R, N
        d.w_f16 = _ptr(w_f16)
        d.bias = _ptr(bias.detach()) if bias is not None else None
        d.skip = _ptr(skip)
        return d

    def conv16(self, x0: torch.Tensor, x1: Optional[torch.Tensor], w_f16: torch.Tensor, N: int, R: int,
               bias: Optional[torch.Tensor], act: int, out: torch.Tensor, aux: Optional[torch.Tensor] = None, aux_kind: int = 0,
               alpha: float = 1.0, alpha_dev: Optional[torch.Tensor] = None, skip: Optional[torch.Tensor] = None,
               amax_out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """odex_conv16(PLAIN): out = act(conv(cat(x0, x1)) + bias), then * act'(aux) (aux_kind 1 / 2) or aux + alpha * v (3).
        amax_out (all three modes): fp32 tensor [B] <- max |sources| per image.  out may be a pair (lo [B,16,16,s], hi [B,16,16,N-s])
        with s a multiple of 64: the output channels land in two tensors (odex_conv16_desc.out_split)."""
        self.require(x0)
        d = self._conv16_desc(x0, x1, w_f16, N, R, bias, skip)
        d.amax_out = _ptr(amax_out)
        d.mode = _lib.C16_PLAIN
        d.act = act
        if aux is not None and aux_kind:
            d.aux, d.ld_aux = _rows(aux)
            d.aux_kind = aux_kind
        d.alpha = float(alpha)
        d.alpha_dev = _ptr(alpha_dev)
        if isinstance(out, (tuple, list)):
            lo, hi = out
            assert lo.shape[3] % 64 == 0 and lo.shape[3] + hi.shape[3] == N and tuple(lo.shape[:3]) == tuple(hi.shape[:3]) == (x0.shape[0], 16, 16)
            d.out, d.ldo = _rows(lo)
            d.out_hi, d.ld_out_hi = _rows(hi)
            d.out_split = lo.shape[3]
        else:
            assert tuple(out.shape) == (x0.shape[0], 16, 16, N)
            d.out, d.ldo = _rows(out)
        _lib.check(self.lib.odex_conv16(C.byref(d), _stream()), "odex_conv16")
        return out

    def conv16_gru_gates(self, x: torch.Tensor, h: torch.Tensor, w_f16: torch.Tensor, R: int, bias, gamma, beta, eps: float,
                         z: torch.Tensor, rh: torch.Tensor, stats: torch.Tensor, g_raw: Optional[torch.Tensor] = None,
                         amax_out: Optional[torch.Tensor] = None) -> None:
        """odex_conv16(GRU_GATES): z, r = sigmoid(GroupNorm(conv(cat(x, h)) + bias)); rh = r * h; g_raw <- the pre-norm conv."""
        self.require(x)
        Cc = h.shape[3]
        d = self._conv16_desc(x, h, w_f16, 2 * Cc, R, bias, None)
        d.mode = _lib.C16_GRU_GATES
        d.amax_out = _ptr(amax_out)
        d.gamma, d.beta, d.eps = gamma.detach().data_ptr(), beta.detach().data_ptr(), float(eps)
        d.h, d.ldh = _rows(h)
        assert z.is_contiguous() and rh.is_contiguous() and (g_raw is None or g_raw.is_contiguous())
        d.z, d.rh, d.stats, d.g_raw = z.data_ptr(), rh.data_ptr(), stats.data_ptr(), _ptr(g_raw)
        _lib.check(self.lib.odex_conv16(C.byref(d), _stream()), "odex_conv16")

    def conv16_gru_out(self, x: torch.Tensor, rh: torch.Tensor, w_f16: torch.Tensor, R: int, bias, gamma, beta, eps: float,
                       z: torch.Tensor, h: torch.Tensor, hnew: torch.Tensor, stats: torch.Tensor,
                       cand: Optional[torch.Tensor] = None, cpre: Optional[torch.Tensor] = None,
                       amax_out: Optional[torch.Tensor] = None) -> None:
        """odex_conv16(GRU_OUT): c = tanh(GroupNorm(conv(cat(x, rh)) + bias)); hnew = (1 - z) * h + z * c."""
        self.require(x)
        Cc = h.shape[3]
        d = self._conv16_desc(x, rh, w_f16, Cc, R, bias, None)
        d.mode = _lib.C16_GRU_OUT
        d.amax_out = _ptr(amax_out)
        d.gamma, d.beta, d.eps = gamma.detach().data_ptr(), beta.detach().data_ptr(), float(eps)
        d.h, d.ldh = _rows(h)
        assert z.is_contiguous() and (cand is None or cand.is_contiguous()) and (cpre is None or cpre.is_contiguous())
        d.z_in = z.data_ptr()
        d.hnew, d.ld_hnew = _rows(hnew)
        d.cand, d.g_raw, d.stats = _ptr(cand), _ptr(cpre), stats.data_ptr()
        _lib.check(self.lib.odex_conv16(C.byref(d), _stream()), "odex_conv16")

    # ------------------------------------------------------------------ decoder tail (LeakyReLU + ConvTranspose2d(C,1,4,2,1))
    @staticmethod
    def decode_tail_supported(p: torch.Tensor) -> bool:
        return p.dim() == 4 and tuple(p.shape[1:]) == (16, 16, 128) and p.shape[0] > 0 and p.is_contiguous()

    def decode_tail_fwd(self, p: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], slope: float,
                        out: torch.Tensor) -> torch.Tensor:
        """odex_decode_tail_fwd: p [n,Hc,Wc,4C] phase-blocked pre-activation -> out [n,4Hc,4Wc]."""
        self.require(p)
        n, hc, wc, c4 = p.shape
        assert p.is_contiguous() and out.is_contiguous() and tuple(out.shape) == (n, 4 * hc, 4 * wc)
        assert w.is_contiguous() and w.numel() == (c4 // 4) * 16
        _lib.check(self.lib.odex_decode_tail_fwd(p.data_ptr(), n, hc, wc, c4 // 4, w.data_ptr(), _ptr(b), float(slope),
                                                 out.data_ptr(), _stream()), "odex_decode_tail_fwd")
        return out

    def decode_tail_bwd(self, p: torch.Tensor, dout: torch.Tensor, w: torch.Tensor, slope: float, dp: torch.Tensor,
                        dw: Optional[torch.Tensor], db: Optional[torch.Tensor]) -> None:
        """odex_decode_tail_bwd: dp = d/dp, dw [C,16] +=, db [1] +=."""
        n, hc, wc, c4 = p.shape
        assert p.is_contiguous() and dout.is_contiguous() and dp.is_contiguous() and dp.shape == p.shape
        assert dout.numel() == n * 16 * hc * wc
        _lib.check(self.lib.odex_decode_tail_bwd(p.data_ptr(), dout.data_ptr(), n, hc, wc, c4 // 4, w.data_ptr(), float(slope),
                                                 dp.data_ptr(), _ptr(dw), _ptr(db), _stream()), "odex_decode_tail_bwd")

    # ------------------------------------------------------------------ encoder stem (Conv2d(1,16,3,2,1) + LeakyReLU -> cells)
    @staticmethod
    def encode_stem_supported(x: torch.Tensor, cout: int) -> bool:
        return x.dim() == 4 and tuple(x.shape[1:]) == (1, 64, 64) and cout == 16 and x.shape[0] > 0

    def encode_stem_fwd(self, x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], slope: float,
                        cells: torch.Tensor) -> torch.Tensor:
        """odex_encode_stem_fwd: x [n,1,64,64] -> cells [n,16,16,64] (space-to-depth of the activated 16-channel output)."""
        self.require(x)
        n = x.shape[0]
        assert x.is_contiguous() and w.is_contiguous() and cells.is_contiguous() and tuple(cells.shape) == (n, 16, 16, 64)
        _lib.check(self.lib.odex_encode_stem_fwd(x.data_ptr(), n, 64, 64, 16, w.data_ptr(), _ptr(b), float(slope),
                                                 cells.data_ptr(), _stream()), "odex_encode_stem_fwd")
        return cells

    def encode_stem_bwd(self, x: torch.Tensor, cells: torch.Tensor, dcells: torch.Tensor, slope: float,
                        dw: Optional[torch.Tensor], db: Optional[torch.Tensor]) -> None:
        """odex_encode_stem_bwd: dw [16,9] +=, db [16] += from the saved cells and their gradient."""
        n = x.shape[0]
        assert x.is_contiguous() and cells.is_contiguous() and dcells.is_contiguous() and dcells.shape == cells.shape
        _lib.check(self.lib.odex_encode_stem_bwd(x.data_ptr(), cells.data_ptr(), dcells.data_ptr(), n, 64, 64, 16, float(slope),
                                                 _ptr(dw), _ptr(db), _stream()), "odex_encode_stem_bwd")

    # ------------------------------------------------------------------ BatchNorm(batch statistics) + ReLU, flow warp (VidODE)
    def bn_relu_fwd(self, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, relu: bool, y: torch.Tensor,
                    stats: torch.Tensor) -> None:
        """odex_bn_relu_fwd on dense NHWC x [..., C]: y = relu(batchnorm(x)); stats [3, C] <- (mean, rstd, biased var)."""
        self.require(x)
        Cc = x.shape[-1]
        M = x.numel() // Cc
        assert x.is_contiguous() and y.is_contiguous() and stats.is_contiguous() and stats.numel() == 3 * Cc
        n = int(self.lib.odex_bn_scratch_floats(M, Cc))
        scratch = torch.empty(n, dtype=torch.float32, device=x.device)
        _lib.check(self.lib.odex_bn_relu_fwd(x.data_ptr(), M, Cc, gamma.detach().data_ptr(), beta.detach().data_ptr(), float(eps),
                                             1 if relu else 0, y.data_ptr(), stats.data_ptr(), scratch.data_ptr(), n, _stream()),
                   "odex_bn_relu_fwd")

    def bn_relu_bwd(self, x, y, dy, gamma, stats, relu: bool, dx, dgb) -> None:
        """odex_bn_relu_bwd: dx and dgb [2, C] <- (dbeta, dgamma)."""
        Cc = x.shape[-1]
        M = x.numel() // Cc
        assert x.is_contiguous() and dy.is_contiguous() and dx.is_contiguous() and dgb.numel() == 2 * Cc
        n = int(self.lib.odex_bn_scratch_floats(M, Cc))
        scratch = torch.empty(n, dtype=torch.float32, device=x.device)
        _lib.check(self.lib.odex_bn_relu_bwd(x.data_ptr(), _ptr(y), dy.data_ptr(), M, Cc, gamma.detach().data_ptr(), stats.data_ptr(),
                                             1 if relu else 0, dx.data_ptr(), dgb.data_ptr(), scratch.data_ptr(), n, _stream()),
                   "odex_bn_relu_bwd")

    def warp_compose_fwd(self, prev: torch.Tensor, head: torch.Tensor, lin_x: torch.Tensor, lin_y: torch.Tensor,
                         warped: torch.Tensor, pred: torch.Tensor) -> None:
        """odex_warp_compose_fwd: prev [B,C,H,W] NCHW, head [B,H,W,>=C+3] NHWC view -> warped, pred [B,C,H,W]."""
        self.require(prev)
        B, Cc, H, W = prev.shape
        ph, ldh = _rows(head)
        assert prev.is_contiguous() and warped.is_contiguous() and pred.is_contiguous() and tuple(head.shape[:3]) == (B, H, W)
        _lib.check(self.lib.odex_warp_compose_fwd(prev.data_ptr(), ph, ldh, lin_x.data_ptr(), lin_y.data_ptr(), B, Cc, H, W,
                                                  warped.data_ptr(), pred.data_ptr(), _stream()), "odex_warp_compose_fwd")

    # ------------------------------------------------------------------ bilinear 2x upsample (NHWC)
    def upsample2x_fwd(self, x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """odex_upsample2x_fwd: x [n,H,W,C] -> out [n,2H,2W,C] (align_corners=False)."""
        self.require(x)
        n, H, W, Cc = x.shape
        assert x.is_contiguous() and out.is_contiguous() and tuple(out.shape) == (n, 2 * H, 2 * W, Cc)
        _lib.check(self.lib.odex_upsample2x_fwd(x.data_ptr(), n, H, W, Cc, out.data_ptr(), _stream()), "odex_upsample2x_fwd")
        return out

    def upsample2x_bwd(self, dout: torch.Tensor, dx: torch.Tensor) -> torch.Tensor:
        """odex_upsample2x_bwd: the adjoint, dout [n,2H,2W,C] -> dx [n,H,W,C]."""
        self.require(dout)
        n, H, W, Cc = dx.shape
        assert dout.is_contiguous() and dx.is_contiguous() and tuple(dout.shape) == (n, 2 * H, 2 * W, Cc)
        _lib.check(self.lib.odex_upsample2x_bwd(dout.data_ptr(), n, H, W, Cc, dx.data_ptr(), _stream()), "odex_upsample2x_bwd")
        return dx

    def conv2d_wgrad(self, x0: torch.Tensor, x1: Optional[torch.Tensor], dy: torch.Tensor, R: int,
                     dw: torch.Tensor, db: Optional[torch.Tensor]) -> None:
        """odex_conv2d_wgrad: dw (OIHW) += X^T dy, db += colsum(dy)."""
        B, H, W, c0 = x0.shape
        p0, ld0 = _rows(x0)
        p1, ld1, c1 = None, 0, 0
        if x1 is not None:
            p1, ld1 = _rows(x1)
            c1 = x1.shape[3]
        pdy, ldy = _rows(dy)
        N = dy.shape[3]
        assert dw.is_contiguous() and tuple(dw.shape) == (N, c0 + c1, R, R)
        _lib.check(self.lib.odex_conv2d_wgrad(p0, ld0, c0, p1, ld1, c1, pdy, ldy, B, H, W, R, N, dw.data_ptr(),
                                              _ptr(db), _stream()), "odex_conv2d_wgrad")

    @staticmethod
    def wgrad_f16_shape_ok(c0: int, c1: int, N: int, H: int, W: int, R: int) -> bool:
        """Mirror of wgrad_f16_supported() in csrc/wgrad_f16.cu."""
        return R in (1, 3, 5) and c0 % 64 == 0 and c1 % 64 == 0 and N % 64 == 0 and W % 16 == 0 and H % 4 == 0 \
            and os.environ.get("ODEX_NO_WGRAD16") is None

    def conv2d_wgrad_packed(self, x0: torch.Tensor, x1: Optional[torch.Tensor], dy: torch.Tensor, R: int,
                            dwp: torch.Tensor, db: Optional[torch.Tensor], impl: int = CONV_AUTO, f16: bool = False,
                            amax: Optional[torch.Tensor] = None) -> None:
        """odex_conv2d_wgrad_packed: dwp [R*R, N, c0+c1] += dy^T X (tensor-core kernel where the shape allows),
        db += colsum(dy).  f16: the kind::f16 kernel with the two-term fp16 split (odex_conv2d_wgrad_f16) where its shape test
        passes.  amax: (ax, ad), fp32 tensors [B] with ax[b] >= max |X[b]|, ad[b] >= max |dy[b]| the caller already holds
        (odex_conv2d_wgrad_f16_scaled: the same kernel without its max pass, operands scaled per image)."""
        B, H, W, c0 = x0.shape
        if (f16 or amax is not None) and impl == CONV_AUTO and self.wgrad_f16_shape_ok(c0, 0 if x1 is None else x1.shape[3], dy.shape[3], H, W, R):
            p0, ld0 = _rows(x0)
            p1, ld1, c1 = None, 0, 0
            if x1 is not None:
                p1, ld1 = _rows(x1)
                c1 = x1.shape[3]
            pdy, ldy = _rows(dy)
            N = dy.shape[3]
            assert dwp.is_contiguous() and tuple(dwp.shape) == (R * R, N, c0 + c1), (tuple(dwp.shape), (R * R, N, c0 + c1))
            if amax is not None:
                ax, ad = amax
                assert ax.dtype == torch.float32 and ad.dtype == torch.float32 and ax.numel() == B and ad.numel() == B
                assert ax.is_contiguous() and ad.is_contiguous()
                _lib.check(self.lib.odex_conv2d_wgrad_f16_scaled(p0, ld0, c0, p1, ld1, c1, pdy, ldy, B, H, W, R, N, dwp.data_ptr(),
                                                                 _ptr(db), ax.data_ptr(), ad.data_ptr(), _stream()),
                           "odex_conv2d_wgrad_f16_scaled")
                return
            scratch = torch.empty(2, dtype=torch.float32, device=x0.device)
            _lib.check(self.lib.odex_conv2d_wgrad_f16(p0, ld0, c0, p1, ld1, c1, pdy, ldy, B, H, W, R, N, dwp.data_ptr(),
                                                      _ptr(db), scratch.data_ptr(), _stream()), "odex_conv2d_wgrad_f16")
            return
        p0, ld0 = _rows(x0)
        p1, ld1, c1 = None, 0, 0
        if x1 is not None:
            p1, ld1 = _rows(x1)
            c1 = x1.shape[3]
        pdy, ldy = _rows(dy)
        N = dy.shape[3]
        assert dwp.is_contiguous() and tuple(dwp.shape) == (R * R, N, c0 + c1), (tuple(dwp.shape), (R * R, N, c0 + c1))
        _lib.check(self.lib.odex_conv2d_wgrad_packed(p0, ld0, c0, p1, ld1, c1, pdy, ldy, B, H, W, R, N, dwp.data_ptr(),
                                                     _ptr(db), impl, _stream()), "odex_conv2d_wgrad_packed")

    def wgrad_unpack(self, dwp: torch.Tensor, dw: torch.Tensor) -> None:
        """odex_wgrad_unpack: dw (OIHW) += dwp ([R*R, N, C])."""
        N, Cc, R, _ = dw.shape
        assert dw.is_contiguous() and dwp.is_contiguous() and tuple(dwp.shape) == (R * R, N, Cc)
        _lib.check(self.lib.odex_wgrad_unpack(dwp.data_ptr(), N, Cc, R, dw.data_ptr(), _stream()), "odex_wgrad_unpack")

    # ------------------------------------------------------------------ ConvGRU gates (+GroupNorm)
    def gru_gates_fwd(self, g, h, gamma, beta, eps, z, rh, stats) -> None:
        B, H, W, C2 = g.shape
        Cc = C2 // 2
        assert g.is_contiguous() and z.is_contiguous() and rh.is_contiguous()
        ph, ldh = _rows(h)
        _lib.check(self.lib.odex_gru_gates_fwd(g.data_ptr(), ph, ldh, gamma.detach().data_ptr(), beta.detach().data_ptr(),
                                               float(eps), B, H * W, Cc, z.data_ptr(), rh.data_ptr(), stats.data_ptr(),
                                               _stream()), "odex_gru_gates_fwd")

    def gru_out_fwd(self, cpre, z, h, gamma, beta, eps, hnew, cand, stats) -> None:
        B, H, W, Cc = cpre.shape
        assert cpre.is_contiguous() and z.is_contiguous() and hnew.is_contiguous()
        ph, ldh = _rows(h)
        _lib.check(self.lib.odex_gru_out_fwd(cpre.data_ptr(), z.data_ptr(), ph, ldh, gamma.detach().data_ptr(),
                                             beta.detach().data_ptr(), float(eps), B, H * W, Cc, hnew.data_ptr(),
                                             _ptr(cand), stats.data_ptr(), _stream()), "odex_gru_out_fwd")

    def gru_out_bwd(self, dhnew, cpre, cand, z, h, gamma, stats, dcpre, dz, dh, dgamma, dbeta) -> None:
        B, H, W, Cc = cpre.shape
        assert dhnew.is_contiguous() and dcpre.is_contiguous() and dz.is_contiguous() and dh.is_contiguous()
        ph, ldh = _rows(h)
        _lib.check(self.lib.odex_gru_out_bwd(dhnew.data_ptr(), cpre.data_ptr(), cand.data_ptr(), z.data_ptr(), ph, ldh,
                                             gamma.detach().data_ptr(), stats.data_ptr(), B, H * W, Cc, dcpre.data_ptr(),
                                             dz.data_ptr(), dh.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                             _stream()), "odex_gru_out_bwd")

    def gru_gates_bwd(self, dz, drh, dh_in, g, h, gamma, beta, stats, dg, dh_out, dgamma, dbeta) -> None:
        B, H, W, C2 = g.shape
        Cc = C2 // 2
        assert dz.is_contiguous() and dh_in.is_contiguous() and dg.is_contiguous()
        pdrh, lddrh = _rows(drh)
        ph, ldh = _rows(h)
        pdo, lddh = _rows(dh_out)
        _lib.check(self.lib.odex_gru_gates_bwd(dz.data_ptr(), pdrh, lddrh, dh_in.data_ptr(), g.data_ptr(), ph, ldh,
                                               gamma.detach().data_ptr(), beta.detach().data_ptr(), stats.data_ptr(),
                                               B, H * W, Cc, dg.data_ptr(), pdo, lddh, dgamma.data_ptr(),
                                               dbeta.data_ptr(), _stream()), "odex_gru_gates_bwd")

    # ------------------------------------------------------------------ Runge-Kutta planes
    def lincomb(self, out: torch.Tensor, ins: Sequence[torch.Tensor], coefs: Sequence[float], dtmask: int = 0,
                dt_dev: Optional[torch.Tensor] = None, skip: Optional[torch.Tensor] = None) -> torch.Tensor:
        """odex_lincomb: out = sum_j c_j * in_j, c_j = coef_j * (dt if bit j of dtmask else 1), j ascending."""
        n = out.numel()
        nin = len(ins)
        assert 1 <= nin <= 8 and len(coefs) == nin
        arr = (C.c_void_p * nin)(*[_flat(t) for t in ins])
        cf = (C.c_double * nin)(*[float(c) for c in coefs])
        for t in ins:
            assert t.numel() == n
        _lib.check(self.lib.odex_lincomb(_flat(out), nin, arr, cf, dtmask, _ptr(dt_dev), n, _ptr(skip), _stream()),
                   "odex_lincomb")
        return out

    def act_bwd(self, out: torch.Tensor, dy: torch.Tensor, a: torch.Tensor, kind: int) -> torch.Tensor:
        """out = dy * act'(a) elementwise (final-activation gradient of a conv stack)."""
        _lib.check(self.lib.odex_act_bwd(_flat(out), _flat(dy), _flat(a), kind, out.numel(), _stream()), "odex_act_bwd")
        return out

    # ------------------------------------------------------------------ dopri5 controller (device resident)
    def dopri_state_bytes(self) -> int:
        return int(self.lib.odex_dopri_state_bytes())

    def dopri_init(self, st, t_dev, n_out, rtol, atol, max_attempts, save_cap=0) -> None:
        _lib.check(self.lib.odex_dopri_init(st.data_ptr(), t_dev.data_ptr(), n_out, rtol, atol, max_attempts, save_cap,
                                            _stream()), "odex_dopri_init")

    def dopri_initnorm(self, st, phase, y0, f0, f1, fuse_ctrl=False) -> None:
        _lib.check(self.lib.odex_dopri_initnorm(st.data_ptr(), phase, _flat(y0), _flat(f0), _ptr(f1), y0.numel(),
                                                1 if fuse_ctrl else 0, _stream()), "odex_dopri_initnorm")

    def dopri_initctrl(self, st, phase) -> None:
        _lib.check(self.lib.odex_dopri_initctrl(st.data_ptr(), phase, _stream()), "odex_dopri_initctrl")

    def dopri_errctrl(self, st, t_dev, y0, y1, k7: List[torch.Tensor], log, log_cap, seq) -> None:
        """K4 fused: error estimate + RMS norm + accept/reject + step-size controller in one launch."""
        arr = (C.c_void_p * 7)(*[_flat(t) for t in k7])
        _lib.check(self.lib.odex_dopri_errctrl(st.data_ptr(), t_dev.data_ptr(), _flat(y0), _flat(y1), arr, y0.numel(),
                                               _ptr(log), log_cap, seq, _stream()), "odex_dopri_errctrl")

    def dopri_errnorm(self, st, y0, y1, k7: List[torch.Tensor]) -> None:
        arr = (C.c_void_p * 7)(*[_flat(t) for t in k7])
        _lib.check(self.lib.odex_dopri_errnorm(st.data_ptr(), _flat(y0), _flat(y1), arr, y0.numel(), _stream()),
                   "odex_dopri_errnorm")

    def dopri_ctrl(self, st, t_dev, log, log_cap, seq) -> None:
        _lib.check(self.lib.odex_dopri_ctrl(st.data_ptr(), t_dev.data_ptr(), _ptr(log), log_cap, seq, _stream()),
                   "odex_dopri_ctrl")

    def dopri_advance(self, st, t_dev, y0, y1, k7: List[torch.Tensor], sol, ysave, fsave, save_cap, seq) -> None:
        arr = (C.c_void_p * 7)(*[_flat(t) for t in k7])
        _lib.check(self.lib.odex_dopri_advance(st.data_ptr(), t_dev.data_ptr(), _flat(y0), _flat(y1), arr, _flat(sol),
                                               y0.numel(), _ptr(ysave), _ptr(fsave), save_cap, seq, _stream()),
                   "odex_dopri_advance")

    def nan_flag(self, x: torch.Tensor, flag: torch.Tensor, bit: int) -> None:
        """odex_nan_flag: *flag |= bit when x holds a NaN (sticky device word, read by the host once per forward)."""
        assert flag.dtype == torch.int32 and x.is_contiguous()
        _lib.check(self.lib.odex_nan_flag(x.data_ptr(), x.numel(), flag.data_ptr(), bit, _stream()), "odex_nan_flag")

    # ------------------------------------------------------------------ layout
    def nchw_to_nhwc(self, x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """x: [..., C, H, W] dense -> out: [..., H, W, C] dense."""
        Cc, H, W = x.shape[-3:]
        Bn = x.numel() // (Cc * H * W)
        assert x.is_contiguous() and out.is_contiguous() and x.dtype == torch.float32
        _lib.check(self.lib.odex_nchw_to_nhwc(x.data_ptr(), out.data_ptr(), Bn, Cc, H * W, _stream()), "odex_nchw_to_nhwc")
        return out

    def nhwc_to_nchw(self, x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        H, W, Cc = x.shape[-3:]
        Bn = x.numel() // (Cc * H * W)
        assert x.is_contiguous() and out.is_contiguous() and x.dtype == torch.float32
        _lib.check(self.lib.odex_nhwc_to_nchw(x.data_ptr(), out.data_ptr(), Bn, Cc, H * W, _stream()), "odex_nhwc_to_nchw")
        return out

    def launch_count(self) -> int:
        return _lib.launch_count()


_backend: Optional[Backend] = None


def backend() -> Backend:
    """The process-wide backend; created on first use (loads libodex.so, raises if it is missing)."""
    global _backend
    if _backend is None:
        _backend = Backend()
    return _backend


def set_backend(b: Optional[Backend]) -> Optional[Backend]:
    """Test hook: swap the backend object (tests/fake_backend.py); returns the previous one."""
    global _backend
    prev, _backend = _backend, b
    return prev
