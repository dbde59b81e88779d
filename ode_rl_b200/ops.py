"""Host-side operators of the hot path, each with an explicit forward and a hand-written backward that
enqueue libodex kernels (through kernels.Backend) on the current stream.  No torch compute ops here:
torch only owns the device buffers.

  ConvNetOp  -- the ODE function `gradient_net` (helpers/utils.py:158-183 via modules/DiffEqSolver.py:71-80)
                and `transform_z0` (modules/ODEConvGRUCell.py:27-30): a chain of stride-1 "same" convolutions
                with ReLU / Tanh in between, optionally fused with `res + alpha * f(x)` (the Euler increment of
                modules/ODEConvGRUCell.py:54-57).
  GruOp      -- one ConvGRU update (modules/ConvGRUCell.py:72-82): conv5x5(cat(x,h)) -> GroupNorm -> sigmoid
                gates -> conv5x5(cat(x, r*h)) -> GroupNorm -> tanh -> blend.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from . import kernels as K
from .kernels import ACT_NONE, ACT_RELU, ACT_TANH, CONV_AUTO, CONV_SIMT, CONV_TC, PackedConv

# "auto": tcgen05 implicit GEMM where the shape allows it, exact-fp32 FFMA kernel otherwise;
# "simt": force the FFMA kernel (cross-check / debugging); "tc": require the tensor-core kernel.
CONV_IMPL = os.environ.get("ODEX_CONV_IMPL", "auto").lower()


# Operand split of the fused conv chain (odex_convnet): "f16" = two-term fp16 split on tcgen05 kind::f16 (22 significant bits
# per operand at twice the tf32 MMA rate), "tf32" = 3xTF32 on kind::tf32.
CONV_SPLIT = os.environ.get("ODEX_SPLIT", "f16").lower()


def set_conv_split(name: str) -> str:
    global CONV_SPLIT
    assert name in ("f16", "tf32")
    prev, CONV_SPLIT = CONV_SPLIT, name
    return prev


def _wgrad_f16() -> bool:
    """The kind::f16 weight-gradient kernel (csrc/wgrad_f16.cu) is opt-in (ODEX_WGRAD_SPLIT=f16): it is parity-green, but its
    per-tensor scales need a max pass over both operands (21 us per batched launch at B = 128) and its fp32 -> fp16 operand
    transform, not the halved MMA time, bounds a tile -- measured 17.4 ms per step against 16.6 ms with the 3xTF32 kernel
    (profiles/r02_step_breakdown_wgrad_f16.log)."""
    return os.environ.get("ODEX_WGRAD_SPLIT", "tf32").lower() == "f16" and CONV_IMPL != "simt"


def _wgrad_scaled() -> bool:
    """Weight gradients of the fused ODE-function chain on the kind::f16 kernel WITHOUT its max pass (odex_conv2d_wgrad_f16_scaled):
    the chain kernels measure max |a_l| / max |d_l| of every layer's operands anyway and leave them in a maxima buffer the caller
    of ConvNetOp.fwd provides (odex_convnet_desc.amax_out).  Default on; ODEX_WGRAD_SCALED=0 keeps the 3xTF32 kernel."""
    return os.environ.get("ODEX_WGRAD_SCALED", "1") != "0" and CONV_IMPL != "simt"


class SavedActs(list):
    """The activations ConvNetOp.fwd(save=True) keeps for the backward, plus the [L, 2, B] maxima buffer their per-image max |a_l|
    went into ([l, 0]; the backward adds max |d_l| in [l, 1]).  None: not measured."""
    maxima: Optional[torch.Tensor] = None


def set_conv_impl(name: str) -> str:
    global CONV_IMPL
    assert name in ("auto", "simt", "tc")
    prev, CONV_IMPL = CONV_IMPL, name
    return prev


def _use_tc(N: int, c0: int, c1: int, H: int, W: int, R: int) -> bool:
    if CONV_IMPL == "simt":
        return False
    ok = K.tc_shape_ok(N, c0, c1, H, W, R)
    if CONV_IMPL == "tc" and not ok:
        raise ValueError("ODEX_CONV_IMPL=tc but shape N=%d c0=%d c1=%d HxW=%dx%d R=%d is not supported by the "
                         "tcgen05 kernel" % (N, c0, c1, H, W, R))
    return ok


def to_nhwc(x: torch.Tensor) -> torch.Tensor:
    """[..., C, H, W] (any strides) -> dense [..., H, W, C]; zero-copy when x already is channels-last."""
    xp = x.movedim(-3, -1)
    if xp.is_contiguous():
        return xp
    be = K.backend()
    if not x.is_contiguous():
        x = x.contiguous()
    out = torch.empty(xp.shape, dtype=x.dtype, device=x.device)
    return be.nchw_to_nhwc(x, out)


def from_nhwc(x: torch.Tensor, dense_nchw: bool = False) -> torch.Tensor:
    """dense [..., H, W, C] -> logical [..., C, H, W].  By default a zero-copy channels-last view (what cuDNN
    prefers for the torch-side ConvDecode); dense_nchw=True materialises NCHW memory with the transpose kernel."""
    if not dense_nchw:
        return x.movedim(-1, -3)
    out = torch.empty(x.movedim(-1, -3).shape, dtype=x.dtype, device=x.device)
    return K.backend().nhwc_to_nchw(x, out)


def live(params):
    """The parameter list an autograd Function of this package is applied to: the parameters themselves when a backward may
    follow, NOTHING under torch.no_grad() -- ctx.needs_input_grad still reports True there for a tensor that requires grad,
    and the Functions key "save activations for the backward" on it (an inference rollout must not keep 36 f-evals of
    activations alive, nor write dopri5's per-step checkpoints)."""
    return tuple(params) if torch.is_grad_enabled() else ()


def live_input(x: torch.Tensor) -> torch.Tensor:
    return x if torch.is_grad_enabled() else x.detach()


class GradSink:
    """fp32 accumulators for parameter gradients, keyed by parameter identity (kernels accumulate with +=)."""

    def __init__(self):
        self._g: Dict[int, torch.Tensor] = {}
        self._packed: Dict[int, torch.Tensor] = {}      # conv weights: [R*R, N, C] accumulators (kernel layout)
        self._pending: Dict[int, Tuple[nn.Conv2d, list]] = {}
        self.deferred = False                           # True: add_wgrad only records; flush() launches, batched

    def add_wgrad(self, conv: nn.Conv2d, x: torch.Tensor, dy: torch.Tensor, amax=None, x1: Optional[torch.Tensor] = None) -> None:
        """dW(conv) += x^T dy, db += colsum(dy).  While `deferred`, the request is only recorded: weight gradients are off
        the critical path of a backward pass, and requests whose operands are consecutive slices of one buffer (the RK
        stages of a step, allocated as [s, B, ...] arenas) run as ONE launch over s*B images.
        amax = (ax, ad): per-image maxima [B] of x and dy, as the chain kernels measured them; such requests run on the kind::f16
        kernel without a max pass, and merge when their maxima are consecutive slices too.  x1: second source of a convolution
        over torch.cat((x, x1)) (the ConvGRU's), merged under the same rule."""
        if (conv.weight is None or not conv.weight.requires_grad) and (conv.bias is None or not conv.bias.requires_grad):
            return
        if not self.deferred:
            self._launch(conv, x, dy, amax, x1)
        else:
            self._pending.setdefault(id(conv.weight), (conv, []))[1].append((x, dy, amax, x1))

    def _launch(self, conv: nn.Conv2d, x: torch.Tensor, dy: torch.Tensor, amax=None, x1: Optional[torch.Tensor] = None) -> None:
        gw, gb = self.get_packed(conv.weight), self.get(conv.bias)
        if gw is None:          # bias-only gradient: the kernel still needs a dw target
            gw = _scratch_packed(conv.weight)
        K.backend().conv2d_wgrad_packed(x, x1, dy, conv.kernel_size[0], gw, gb, f16=_wgrad_f16(), amax=amax)

    def flush(self) -> None:
        for conv, items in self._pending.values():
            # requests with measured maxima merge among themselves (one kind::f16 launch per run), the rest likewise
            for group in ([it for it in items if it[2] is not None], [it for it in items if it[2] is None]):
                for run in _slice_runs(group):
                    self._launch(conv, *run)
        self._pending.clear()

    def get_packed(self, w: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """Accumulator of a Conv2d weight gradient in the wgrad kernels' layout; converted to OIHW in result()."""
        if w is None or not w.requires_grad:
            return None
        g = self._packed.get(id(w))
        if g is None:
            N, Cc, R, _ = w.shape
            g = torch.zeros((R * R, N, Cc), dtype=torch.float32, device=w.device)
            self._packed[id(w)] = g
        return g

    def get(self, p: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if p is None or not p.requires_grad:
            return None
        g = self._g.get(id(p))
        if g is None:
            g = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
            self._g[id(p)] = g
        return g

    def result(self, params: Sequence[torch.Tensor]) -> Tuple[Optional[torch.Tensor], ...]:
        be = K.backend()
        self.flush()
        for p in params:
            gp = self._packed.pop(id(p), None)
            if gp is not None:
                be.wgrad_unpack(gp, self.get(p))
        return tuple(self._g.get(id(p)) if p.requires_grad else None for p in params)


def _slice_runs(items):
    """[(x, dy[, amax[, x1]])] -> the same requests with every run of consecutive dense slices of one storage merged into one
    (x [n*B, ...], dy [n*B, ...], amax, x1) request.  amax = (ax, ad) per-image maxima [B] or None, x1 = second source or None;
    a run merges only while EVERY tensor of the request (x, dy, and when present x1, ax, ad) continues its predecessor's, and
    comes back with all of them n*B long.  Requests given as (x, dy) pairs come back as pairs."""
    pairs = all(len(it) == 2 for it in items)
    items = [(it[0], it[1], it[2] if len(it) > 2 else None, it[3] if len(it) > 3 else None) for it in items]
    items = sorted(items, key=lambda it: it[0].data_ptr())

    def parts(it):
        x, dy, amax, x1 = it
        return [x, dy] + ([x1] if x1 is not None else []) + (list(amax) if amax is not None else [])

    def follows(t0, t1, n):
        return (t1.shape == t0.shape and t1.stride() == t0.stride()
                and t1.data_ptr() == t0.data_ptr() + n * t0.shape[0] * t0.stride(0) * t0.element_size()
                and t1.untyped_storage().data_ptr() == t0.untyped_storage().data_ptr())

    def mergeable(t):
        # dense image slices: [B, H, W, C] whose rows may be a channel slice of a wider buffer, images back to back
        if t.dim() == 1:
            return t.is_contiguous()
        return t.dim() == 4 and t.stride(3) == 1 and t.stride(1) == t.shape[2] * t.stride(2) and t.stride(0) == t.shape[1] * t.stride(1)

    out, i = [], 0
    while i < len(items):
        p0 = parts(items[i])
        n = 1
        ok = all(mergeable(t) for t in p0)
        while ok and i + n < len(items):
            it1 = items[i + n]
            if (it1[2] is None) != (items[i][2] is None) or (it1[3] is None) != (items[i][3] is None):
                break
            if not all(follows(t0, t1, n) for t0, t1 in zip(p0, parts(it1))):
                break
            n += 1
        if n > 1:
            m = [torch.as_strided(t, (n * t.shape[0],) + tuple(t.shape[1:]), t.stride()) for t in p0]
            x, dy = m[0], m[1]
            k = 2
            x1 = None
            if items[i][3] is not None:
                x1 = m[k]
                k += 1
            amax = (m[k], m[k + 1]) if items[i][2] is not None else None
            run = (x, dy, amax, x1)
        else:
            run = items[i]
        out.append(run[:2] if pairs else run)
        i += n
    return out


def _scratch_packed(w: torch.Tensor) -> torch.Tensor:
    N, Cc, R, _ = w.shape
    return torch.zeros((R * R, N, Cc), dtype=torch.float32, device=w.device)


class _PackCache:
    """Kernel-layout copies of a Conv2d weight, refreshed when the parameter changes (optimizer step / load)."""

    def __init__(self, conv: nn.Conv2d):
        self.conv = conv
        self.key = None
        self.pk: Optional[PackedConv] = None
        self.shape_key = None
        self.key16 = None
        self.pk16: Optional[PackedConv] = None

    def get_f16(self, need_dgrad: bool) -> PackedConv:
        """The two-term fp16 split copies (odex_pack_weights tc = 2) the fused chain reads when CONV_SPLIT == 'f16'."""
        w = self.conv.weight
        key = (w.data_ptr(), w._version, w.device)
        if self.pk16 is None or key != self.key16:
            self.pk16 = PackedConv(w.shape[0], w.shape[1], w.shape[2])
            self.key16 = key
        pk = self.pk16
        be = K.backend()
        if need_dgrad and pk.fwd_f16 is None and pk.dgrad_f16 is None and hasattr(be, "pack_weights_f16_pair") \
                and w.shape[0] % 64 == 0 and w.shape[1] % 64 == 0 and w.is_contiguous():
            pk.fwd_f16, pk.dgrad_f16 = be.pack_weights_f16_pair(w)      # a training step needs both: one pair of launches
        if pk.fwd_f16 is None:
            pk.fwd_f16 = be.pack_weights(w, dgrad=False, tc=2)
        if need_dgrad and pk.dgrad_f16 is None:
            pk.dgrad_f16 = be.pack_weights(w, dgrad=True, tc=2)
        return pk

    def get(self, H: int, W: int, c0: int, c1: int, need_dgrad: bool) -> PackedConv:
        w = self.conv.weight
        key = (w.data_ptr(), w._version, w.device, CONV_IMPL)
        skey = (H, W, c0, c1)
        if self.pk is None or key != self.key or skey != self.shape_key:
            self.pk = PackedConv(w.shape[0], w.shape[1], w.shape[2])
            self.key, self.shape_key = key, skey
        pk = self.pk
        be = K.backend()
        N, Cin, R = pk.N, pk.C, pk.R
        if _use_tc(N, c0, c1, H, W, R):
            if pk.fwd_tc is None:
                pk.fwd_tc = be.pack_weights(w, dgrad=False, tc=True)
        elif pk.fwd_simt is None:
            pk.fwd_simt = be.pack_weights(w, dgrad=False, tc=False)
        if need_dgrad:
            # the data-gradient operator maps N channels of dy to Cin channels (one source)
            if _use_tc(Cin, N, 0, H, W, R):
                if pk.dgrad_tc is None:
                    pk.dgrad_tc = be.pack_weights(w, dgrad=True, tc=True)
            elif pk.dgrad_simt is None:
                pk.dgrad_simt = be.pack_weights(w, dgrad=True, tc=False)
        return pk

    def invalidate(self):
        self.key = None
        self.pk = None
        self.key16 = None
        self.pk16 = None


def _check_conv(conv: nn.Conv2d, what: str) -> int:
    k = conv.kernel_size
    if not (isinstance(conv, nn.Conv2d) and k[0] == k[1] and k[0] in (1, 3, 5) and conv.stride == (1, 1)
            and conv.padding == ((k[0] - 1) // 2,) * 2 and conv.dilation == (1, 1) and conv.groups == 1
            and conv.padding_mode == "zeros"):
        raise NotImplementedError(
            "%s: the CUDA path takes stride-1 'same' Conv2d with a square 1/3/5 kernel only (got %r)" % (what, conv))
    return k[0]


def _act_kind(m: nn.Module) -> int:
    if isinstance(m, nn.ReLU):
        return ACT_RELU
    if isinstance(m, nn.Tanh):
        return ACT_TANH
    raise NotImplementedError("activation %r is not supported by the CUDA path (ReLU / Tanh only)" % (m,))


def parse_sequential(seq: nn.Sequential, what: str) -> List[Tuple[nn.Conv2d, int]]:
    """[Conv2d, act, Conv2d, act, ..., Conv2d(, act)] -> [(conv, act_after)]."""
    layers: List[Tuple[nn.Conv2d, int]] = []
    mods = list(seq)
    i = 0
    while i < len(mods):
        conv = mods[i]
        if not isinstance(conv, nn.Conv2d):
            raise NotImplementedError("%s: expected Conv2d at position %d of %r" % (what, i, seq))
        _check_conv(conv, what)
        act = ACT_NONE
        if i + 1 < len(mods) and not isinstance(mods[i + 1], nn.Conv2d):
            act = _act_kind(mods[i + 1])
            i += 1
        layers.append((conv, act))
        i += 1
    if not layers:
        raise NotImplementedError("%s: empty conv stack" % what)
    return layers


class ConvNetOp:
    """f(x) = conv_{L-1}(act(... act(conv_0(x)))) (+ trailing act), NHWC in / NHWC out."""

    def __init__(self, seq: nn.Sequential, what: str = "gradient_net"):
        self.layers = parse_sequential(seq, what)
        self.caches = [_PackCache(c) for c, _ in self.layers]
        self.c_in = self.layers[0][0].in_channels
        self.c_out = self.layers[-1][0].out_channels

    def maxima_ok(self, x: torch.Tensor) -> bool:
        """True when the chain kernels can measure the operand maxima of this input's weight gradients (see new_maxima)."""
        if CONV_SPLIT != "f16" or not _wgrad_scaled() or not self._fused(K.backend(), x):
            return False
        return K.backend().wgrad_f16_shape_ok(64, 0, 64, x.shape[1], x.shape[2], 3)

    def new_maxima(self, x: torch.Tensor) -> Optional[torch.Tensor]:
        """An [L, 2, B] maxima buffer for fwd(save=True, maxima=...) when the weight gradients of this chain can use it (fused
        kind::f16 chain + odex_conv2d_wgrad_f16_scaled), else None.  [l, 0, b] <- max |a_l[b]| (forward launch), [l, 1, b] <-
        max |d_l[b]| (backward launch): per image, so launches of any magnitude may later share one weight-gradient launch."""
        if not self.maxima_ok(x):
            return None
        return torch.empty((len(self.layers), 2, x.shape[0]), dtype=torch.float32, device=x.device)

    def params(self) -> List[torch.Tensor]:
        ps = []
        for conv, _ in self.layers:
            ps.append(conv.weight)
            if conv.bias is not None:
                ps.append(conv.bias)
        return ps

    def flops(self, B: int, H: int, W: int) -> int:
        return sum(2 * B * H * W * c.out_channels * c.in_channels * c.kernel_size[0] ** 2 for c, _ in self.layers)

    def invalidate(self):
        for c in self.caches:
            c.invalidate()

    def _fused(self, be, x: torch.Tensor) -> bool:
        """True when the whole chain runs as one odex_convnet launch (16x16 images, every layer 64 -> 64 3x3, tcgen05)."""
        if CONV_IMPL == "simt" or not hasattr(be, "convnet") or not be.convnet_supported(x, len(self.layers)):
            return False
        return all(c.in_channels == 64 and c.out_channels == 64 and c.kernel_size == (3, 3) for c, _ in self.layers)

    # out = res + alpha * f(x)   (res None: out = f(x))
    def fwd(self, x: torch.Tensor, save: bool = False, out: Optional[torch.Tensor] = None,
            res: Optional[torch.Tensor] = None, alpha: float = 1.0, alpha_dev: Optional[torch.Tensor] = None,
            skip: Optional[torch.Tensor] = None, need_dgrad: bool = False,
            bufs: Optional[List[torch.Tensor]] = None, maxima: Optional[torch.Tensor] = None):
        """bufs: optional caller-owned buffers for the outputs of layers 0 .. L-2 (the saved activations).  maxima: a buffer from
        new_maxima() (or an [L, 2, B] view with dense rows); the chain writes max |a_l| per image and the saved activations
        carry it to bwd()."""
        be = K.backend()
        B, H, W, _ = x.shape
        L = len(self.layers)
        saved = SavedActs([x]) if save else None
        a = x
        if self._fused(be, x):
            f16 = CONV_SPLIT == "f16"
            amax_out = None
            if f16 and save and maxima is not None:
                amax_out = [maxima[l, 0] for l in range(L)]
                saved.maxima = maxima
            if f16:
                wts = [self.caches[l].get_f16(save or need_dgrad).fwd_f16 for l in range(L)]
            else:
                wts = [self.caches[l].get(H, W, 64, 0, need_dgrad or save).fwd_tc for l in range(L)]
            last_act = self.layers[-1][1]
            fuse = res is not None and last_act == ACT_NONE
            outs: List[Optional[torch.Tensor]] = [None] * L
            for l in range(L - 1):
                if save:
                    outs[l] = bufs[l] if bufs is not None else torch.empty((B, H, W, 64), dtype=torch.float32, device=x.device)
            if out is not None and (res is None or fuse):
                outs[L - 1] = out
            else:
                outs[L - 1] = torch.empty((B, H, W, 64), dtype=torch.float32, device=x.device)
            aux: List[Optional[torch.Tensor]] = [None] * L
            kinds = [0] * L
            if fuse:
                aux[L - 1], kinds[L - 1] = res, 3
            be.convnet(x, wts, [c.bias for c, _ in self.layers], [act for _, act in self.layers],
                       aux, kinds, outs, alpha=alpha if fuse else 1.0, alpha_dev=alpha_dev if fuse else None, skip=skip,
                       split=1 if f16 else 0, **({"amax_out": amax_out} if amax_out is not None else {}))
            if save:
                saved.extend(outs[:L - 1])
                if last_act != ACT_NONE:
                    saved.append(outs[L - 1])
            a = outs[L - 1]
            if res is not None and not fuse:
                o = out if out is not None else torch.empty_like(a)
                if alpha_dev is not None:
                    be.lincomb(o, [res, a], [1.0, alpha], dtmask=2, dt_dev=alpha_dev, skip=skip)
                else:
                    be.lincomb(o, [res, a], [1.0, alpha], skip=skip)
                a = o
            return (a, saved) if save else a
        for l, (conv, act) in enumerate(self.layers):
            last = l == L - 1
            pk = self.caches[l].get(H, W, a.shape[3], 0, need_dgrad or save)
            fuse = last and (res is not None) and act == ACT_NONE
            if last and out is not None and (res is None or fuse):
                o = out
            elif not last and bufs is not None:
                o = bufs[l]
            else:
                o = torch.empty((B, H, W, pk.N), dtype=torch.float32, device=x.device)
            be.conv2d(a, None, pk.fwd_simt, pk.fwd_tc, pk.N, pk.R, conv.bias, act, o,
                      res=res if fuse else None, alpha=alpha if fuse else 1.0, alpha_dev=alpha_dev if fuse else None,
                      skip=skip)
            if save and (not last or act != ACT_NONE):
                saved.append(o)
            a = o
        if res is not None and self.layers[-1][1] != ACT_NONE:
            # trailing activation: the residual cannot ride in the conv epilogue (the saved activation is needed)
            o = out if out is not None else torch.empty_like(a)
            if alpha_dev is not None:
                be.lincomb(o, [res, a], [1.0, alpha], dtmask=2, dt_dev=alpha_dev, skip=skip)
            else:
                be.lincomb(o, [res, a], [1.0, alpha], skip=skip)
            a = o
        return (a, saved) if save else a

    # dx = res + J_f(x)^T dy ; parameter gradients accumulated into `sink`
    def bwd(self, saved: List[torch.Tensor], dy: torch.Tensor, sink: GradSink, out: Optional[torch.Tensor] = None,
            res: Optional[torch.Tensor] = None, need_dx: bool = True,
            gbufs: Optional[List[torch.Tensor]] = None) -> Optional[torch.Tensor]:
        """gbufs: optional caller-owned buffers for the gradients w.r.t. the inputs of layers 1 .. L-1 (index l-1)."""
        be = K.backend()
        x = saved[0]
        B, H, W, _ = x.shape
        L = len(self.layers)
        d = dy
        last_act = self.layers[-1][1]
        if last_act != ACT_NONE:
            assert d.is_contiguous() and saved[L].is_contiguous()
            d = be.act_bwd(torch.empty_like(saved[L]), d, saved[L], last_act)
        if self._fused(be, x) and d.shape[3] == 64 and (need_dx or L > 1):
            # the data-gradient chain d_{l-1} = conv_l^T(d_l) * act'(a_l) in one launch, then the weight gradients
            lo = 0 if need_dx else 1                 # lowest layer whose input gradient is needed
            ls = list(range(L - 1, lo - 1, -1))      # chain position k handles layer ls[k]
            f16 = CONV_SPLIT == "f16"
            if f16:
                wts = [self.caches[l].get_f16(True).dgrad_f16 for l in ls]
            else:
                wts = [self.caches[l].get(H, W, 64, 0, True).dgrad_tc for l in ls]
            aux: List[Optional[torch.Tensor]] = []
            kinds = []
            outs: List[Optional[torch.Tensor]] = []
            for l in ls:
                if l > 0:
                    aux.append(saved[l]); kinds.append(self.layers[l - 1][1])
                    outs.append(gbufs[l - 1] if gbufs is not None else
                                torch.empty((B, H, W, 64), dtype=torch.float32, device=x.device))
                else:
                    aux.append(res); kinds.append(3 if res is not None else 0)
                    outs.append(out if out is not None else torch.empty((B, H, W, 64), dtype=torch.float32, device=x.device))
            # chain position k reads d_{ls[k]}: its measured max is the dy bound of layer ls[k]'s weight gradient (d_{lo-1}, the
            # chain's last output, is read by no position: with need_dx False layer 0's weight gradient has no bound)
            am = getattr(saved, "maxima", None) if f16 else None
            extra = {}
            if am is not None:
                extra["amax_out"] = [am[l, 1] for l in ls]
            be.convnet(d, wts, [None] * len(ls), [ACT_NONE] * len(ls), aux, kinds, outs, alpha=1.0, split=1 if f16 else 0, **extra)
            grads = {L - 1: d}
            for k, l in enumerate(ls):
                grads[l - 1] = outs[k]               # outs[k] = gradient w.r.t. the input of layer l = d_{l-1}
            for l in range(L - 1, -1, -1):
                sink.add_wgrad(self.layers[l][0], saved[l], grads[l],
                               amax=(am[l, 0], am[l, 1]) if am is not None and l >= lo else None)
            return grads[-1] if need_dx else None
        for l in range(L - 1, -1, -1):
            conv, _ = self.layers[l]
            a_l = saved[l]
            sink.add_wgrad(conv, a_l, d)
            if l == 0 and not need_dx:
                return None
            pk = self.caches[l].get(H, W, a_l.shape[3], 0, True)
            if l > 0:
                nd = gbufs[l - 1] if gbufs is not None else torch.empty((B, H, W, pk.C), dtype=torch.float32, device=x.device)
                be.conv2d(d, None, pk.dgrad_simt, pk.dgrad_tc, pk.C, pk.R, None, ACT_NONE, nd,
                          dact=a_l, dact_kind=self.layers[l - 1][1])
                d = nd
            else:
                dx = out if out is not None else torch.empty((B, H, W, pk.C), dtype=torch.float32, device=x.device)
                be.conv2d(d, None, pk.dgrad_simt, pk.dgrad_tc, pk.C, pk.R, None, ACT_NONE, dx, res=res, alpha=1.0)
                return dx
        return None


class GruOp:
    """One ConvGRU update with GroupNorm gates (modules/ConvGRUCell.py:72-82; update gate first, reset second)."""

    def __init__(self, conv_gates: nn.Sequential, conv_can: nn.Sequential, hidden_dim: int, input_dim: int):
        self.cg, self.gn_g = conv_gates[0], conv_gates[1]
        self.cc, self.gn_c = conv_can[0], conv_can[1]
        self.C, self.Cx = hidden_dim, input_dim
        self.R = _check_conv(self.cg, "ConvGRUCell.conv_gates")
        if _check_conv(self.cc, "ConvGRUCell.conv_can") != self.R:
            raise NotImplementedError("ConvGRUCell: conv_gates and conv_can must share the kernel size")
        for gn, ch in ((self.gn_g, 2 * hidden_dim), (self.gn_c, hidden_dim)):
            if not isinstance(gn, nn.GroupNorm) or gn.num_channels != ch or ch % 32 != 0 or gn.num_groups * 32 != ch \
                    or not gn.affine:
                raise NotImplementedError("ConvGRUCell: the CUDA path needs affine GroupNorm with 32-channel groups "
                                          "(got %r for %d channels)" % (gn, ch))
        if self.cg.bias is None or self.cc.bias is None:
            raise NotImplementedError("ConvGRUCell: bias=False is not supported by the CUDA path")
        self.pc_g, self.pc_c = _PackCache(self.cg), _PackCache(self.cc)

    def params(self) -> List[torch.Tensor]:
        return [self.cg.weight, self.cg.bias, self.gn_g.weight, self.gn_g.bias,
                self.cc.weight, self.cc.bias, self.gn_c.weight, self.gn_c.bias]

    def flops(self, B: int, H: int, W: int) -> int:
        return 2 * B * H * W * (self.Cx + self.C) * self.R * self.R * 3 * self.C

    def _fused16(self, be, x: torch.Tensor, h: torch.Tensor) -> bool:
        """True when the step runs on odex_conv16 (16x16 latents, 64-channel input and hidden state, 3x3 / 5x5, f16 split)."""
        return (CONV_SPLIT == "f16" and CONV_IMPL != "simt" and hasattr(be, "conv16") and self.C == 64 and self.Cx == 64
                and be.conv16_shape_ok(x, h, 2 * self.C, self.R))

    def fwd(self, x: torch.Tensor, h: torch.Tensor, save: bool = False, out: Optional[torch.Tensor] = None,
            bufs: Optional[dict] = None):
        """bufs (fused path, optional): caller-owned {"rh": [B,H,W,C], "mx": [2,2,B]} -- slices of arenas over the steps of a
        recurrence, so that the steps' weight gradients can run as one launch (GradSink.add_wgrad)."""
        be = K.backend()
        B, H, W, Cx = x.shape
        C, R = self.C, self.R
        dev = x.device
        f32 = torch.float32
        if self._fused16(be, x, h):
            # K2 fused (north_star (c)): each convolution and the gate math behind it in ONE launch -- GroupNorm statistics,
            # sigmoid / r*h and tanh / blend ride in the epilogue of the kind::f16 image-resident convolution
            wg = self.pc_g.get_f16(save).fwd_f16
            wc = self.pc_c.get_f16(save).fwd_f16
            z = torch.empty((B, H, W, C), dtype=f32, device=dev)
            rh = bufs["rh"] if bufs is not None else torch.empty((B, H, W, C), dtype=f32, device=dev)
            st_g = torch.empty((B * (2 * C // 32) * 2,), dtype=f32, device=dev)
            st_c = torch.empty((B * (C // 32) * 2,), dtype=f32, device=dev)
            g = torch.empty((B, H, W, 2 * C), dtype=f32, device=dev) if save else None
            # [2, 2, B] per-image maxima of this step: [0] = conv_gates {max |(x, h)|, max |dg|}, [1] = conv_can {max |(x, rh)|,
            # max |dcpre|}; the conv16 launches measure them for their operand scales anyway, the weight gradients take them
            mx = None
            if save and _wgrad_scaled() and be.wgrad_f16_shape_ok(Cx, C, 2 * C, H, W, R):
                mx = bufs["mx"] if bufs is not None and bufs.get("mx") is not None else torch.empty((2, 2, B), dtype=f32, device=dev)
            mxa = (lambda i, j: {"amax_out": mx[i, j]}) if mx is not None else (lambda i, j: {})
            be.conv16_gru_gates(x, h, wg, R, self.cg.bias, self.gn_g.weight, self.gn_g.bias, self.gn_g.eps, z, rh, st_g, g_raw=g,
                                **mxa(0, 0))
            hnew = out if out is not None else torch.empty((B, H, W, C), dtype=f32, device=dev)
            cpre = torch.empty((B, H, W, C), dtype=f32, device=dev) if save else None
            cand = torch.empty((B, H, W, C), dtype=f32, device=dev) if save else None
            be.conv16_gru_out(x, rh, wc, R, self.cc.bias, self.gn_c.weight, self.gn_c.bias, self.gn_c.eps, z, h, hnew, st_c,
                              cand=cand, cpre=cpre, **mxa(1, 0))
            if save:
                return hnew, (x, h, g, z, rh, st_g, cpre, cand, st_c, mx)
            return hnew
        pg = self.pc_g.get(H, W, Cx, C, save)
        g = torch.empty((B, H, W, 2 * C), dtype=f32, device=dev)
        be.conv2d(x, h, pg.fwd_simt, pg.fwd_tc, 2 * C, R, self.cg.bias, ACT_NONE, g)
        z = torch.empty((B, H, W, C), dtype=f32, device=dev)
        rh = torch.empty((B, H, W, C), dtype=f32, device=dev)
        st_g = torch.empty((B * (2 * C // 32) * 2,), dtype=f32, device=dev)
        be.gru_gates_fwd(g, h, self.gn_g.weight, self.gn_g.bias, self.gn_g.eps, z, rh, st_g)
        pc = self.pc_c.get(H, W, Cx, C, save)
        cpre = torch.empty((B, H, W, C), dtype=f32, device=dev)
        be.conv2d(x, rh, pc.fwd_simt, pc.fwd_tc, C, R, self.cc.bias, ACT_NONE, cpre)
        hnew = out if out is not None else torch.empty((B, H, W, C), dtype=f32, device=dev)
        cand = torch.empty((B, H, W, C), dtype=f32, device=dev) if save else None
        st_c = torch.empty((B * (C // 32) * 2,), dtype=f32, device=dev)
        be.gru_out_fwd(cpre, z, h, self.gn_c.weight, self.gn_c.bias, self.gn_c.eps, hnew, cand, st_c)
        if save:
            return hnew, (x, h, g, z, rh, st_g, cpre, cand, st_c)
        return hnew

    def bwd(self, saved, dhnew: torch.Tensor, sink: GradSink, out: Optional[torch.Tensor] = None,
            bufs: Optional[dict] = None, split_out=None):
        """Returns D = [dx | dh] as one [B,H,W,Cx+C] buffer (channel slices are the two gradients).  bufs (optional): caller-owned
        {"dcpre": [B,H,W,C], "dg": [B,H,W,2C]} arena slices (see fwd).  split_out = (dx [B,H,W,Cx], dh [B,H,W,C]) (fused path
        only): the two gradients land in these two DENSE tensors and the pair is returned instead of D."""
        be = K.backend()
        x, h, g, z, rh, st_g, cpre, cand, st_c = saved[:9]
        mx = saved[9] if len(saved) > 9 else None       # the step's maxima (fused forward only)
        B, H, W, Cx = x.shape
        C, R = self.C, self.R
        dev, f32 = x.device, torch.float32

        def grad_or_scratch(p):
            t = sink.get(p)
            return t if t is not None else torch.zeros_like(p)

        dcpre = bufs["dcpre"] if bufs is not None else torch.empty((B, H, W, C), dtype=f32, device=dev)
        dz = torch.empty((B, H, W, C), dtype=f32, device=dev)
        dh_a = torch.empty((B, H, W, C), dtype=f32, device=dev)
        be.gru_out_bwd(dhnew, cpre, cand, z, h, self.gn_c.weight, st_c, dcpre, dz, dh_a,
                       grad_or_scratch(self.gn_c.weight), grad_or_scratch(self.gn_c.bias))
        f16 = self._fused16(be, x, h)
        mx = mx if f16 else None
        D1 = torch.empty((B, H, W, Cx + C), dtype=f32, device=dev)            # [dx_can | drh]
        if f16:
            # the data-gradient launch reads dcpre: its max |dcpre| completes row 1 of the maxima before the weight gradient runs
            be.conv16(dcpre, None, self.pc_c.get_f16(True).dgrad_f16, Cx + C, R, None, ACT_NONE, D1,
                      **({"amax_out": mx[1, 1]} if mx is not None else {}))
        else:
            pc = self.pc_c.get(H, W, Cx, C, True)
            be.conv2d(dcpre, None, pc.dgrad_simt, pc.dgrad_tc, Cx + C, R, None, ACT_NONE, D1)
        sink.add_wgrad(self.cc, x, dcpre, amax=(mx[1, 0], mx[1, 1]) if mx is not None else None, x1=rh)
        dg = bufs["dg"] if bufs is not None else torch.empty((B, H, W, 2 * C), dtype=f32, device=dev)
        drh = D1[..., Cx:]
        be.gru_gates_bwd(dz, drh, dh_a, g, h, self.gn_g.weight, self.gn_g.bias, st_g, dg, drh,
                         grad_or_scratch(self.gn_g.weight), grad_or_scratch(self.gn_g.bias))   # D1[..., Cx:] <- dh_a + drh*r
        if split_out is not None:
            assert f16 and Cx % 64 == 0, "split_out needs the odex_conv16 path"
            D2 = tuple(split_out)
        else:
            D2 = out if out is not None else torch.empty((B, H, W, Cx + C), dtype=f32, device=dev)
        if f16:
            pgf = self.pc_g.get_f16(True).dgrad_f16
            be.conv16(dg[..., :C], dg[..., C:], pgf, Cx + C, R, None, ACT_NONE, D2, aux=D1, aux_kind=3, alpha=1.0,
                      **({"amax_out": mx[0, 1]} if mx is not None else {}))
        else:
            pg = self.pc_g.get(H, W, Cx, C, True)
            be.conv2d(dg, None, pg.dgrad_simt, pg.dgrad_tc, Cx + C, R, None, ACT_NONE, D2, res=D1, alpha=1.0)
        sink.add_wgrad(self.cg, x, dg, amax=(mx[0, 0], mx[0, 1]) if mx is not None else None, x1=h)
        return D2
