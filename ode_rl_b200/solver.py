"""odeint for the latent Neural-ODE rollout: fixed-grid euler / rk4 (3/8 rule) and adaptive dopri5, with
torchdiffeq 0.2.1 semantics (the third-party solver modules/DiffEqSolver.py:9,45-46 calls),
driven from Python but computed entirely by libodex kernels.  Host code only enqueues: the step size, the
accept/reject decision and the output bookkeeping of dopri5 live in device memory (odex_dopri_state) and the host
reads that state back once per forward (per speculative chunk when the guess was too short).

The backward pass is the exact gradient of the discrete computation (the reference uses plain `odeint`, not the
adjoint method -- SURVEY.md fact 5): it replays the saved stages in reverse with the same conv kernels
(data-gradient = forward kernel on flipped weights) and the same stage-combination kernel (transposed tableau).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from . import _lib
from . import kernels as K
from .ops import ConvNetOp, GradSink, from_nhwc, live, live_input, to_nhwc


class Tableau:
    """Explicit Runge-Kutta tableau: stage inputs y_i = y + dt * sum_j beta[i-1][j] k_j, y' = y + dt * sum c_sol[i] k_i."""

    def __init__(self, name: str, beta: Sequence[Sequence[float]], c_sol: Sequence[float]):
        self.name = name
        self.beta = [list(r) for r in beta]
        self.c_sol = list(c_sol)
        self.s = len(c_sol)
        assert len(self.beta) == self.s - 1


EULER = Tableau("euler", [], [1.0])
# torchdiffeq fixed_grid.RK4 = rk_common.rk4_alt_step_func, the 3/8 rule -- not classical RK4
RK4_38 = Tableau("rk4", [[1.0 / 3.0], [-1.0 / 3.0, 1.0], [1.0, -1.0, 1.0]], [0.125, 0.375, 0.375, 0.125])

DP_BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
DP_C_MID = [6025192743 / 30085553152 / 2, 0.0, 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
            187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2]

FIXED = {"euler": EULER, "rk4": RK4_38}


def _lin(be, out, ins, coefs, dt_dev=None, dt_from=0, skip=None):
    """out = sum coefs[j]*ins[j], dropping zero coefficients; inputs j >= dt_from are scaled by *dt_dev."""
    keep = [(t, c, j >= dt_from) for j, (t, c) in enumerate(zip(ins, coefs)) if c != 0.0]
    assert keep, "lincomb with all-zero coefficients"
    mask = 0
    if dt_dev is not None:
        for j, (_, _, m) in enumerate(keep):
            if m:
                mask |= 1 << j
    return be.lincomb(out, [t for t, _, _ in keep], [c for _, c, _ in keep], dtmask=mask,
                      dt_dev=dt_dev if mask else None, skip=skip)


def rk_step_fwd(be, op: ConvNetOp, tab: Tableau, y: torch.Tensor, y_next: torch.Tensor, dt_dev: torch.Tensor,
                save: bool, skip=None, k_bufs: Optional[List[torch.Tensor]] = None, k0: Optional[torch.Tensor] = None,
                step_arenas=None):
    """One explicit RK step y -> y_next with step size *dt_dev (fp64 on the device).  Returns the per-stage saved
    activations (for the backward) when save=True.  k_bufs: s reusable planes for the stage derivatives.  step_arenas:
    (yi [s, B, ...], [per layer [s, B, ...]], maxima [L, 2, s, B] or None) views of the caller's rollout-wide arenas
    (RolloutArenas) instead of per-step allocations."""
    s = tab.s
    saved = []
    if s == 1 and k0 is None:
        # euler: y' = y + dt f(y) rides in the last conv's epilogue (res + alpha*conv)
        r = op.fwd(y, save=save, out=y_next, res=y, alpha=1.0, alpha_dev=dt_dev, skip=skip)
        return [r[1]] if save else None
    ks: List[torch.Tensor] = []
    # training: stage inputs and saved activations of the s stages live in [s, B, ...] arenas, so that the deferred weight
    # gradients of a layer run as one launch over all stages (GradSink.add_wgrad)
    arenas = yi_arena = None
    if step_arenas is not None:
        yi_arena, arenas = step_arenas[0], step_arenas[1]
    elif save and len(op.layers) > 1:
        yi_arena = torch.empty((s,) + tuple(y.shape), dtype=y.dtype, device=y.device)
        if all(c.out_channels == y.shape[3] for c, _ in op.layers[:-1]):
            arenas = [torch.empty((s,) + tuple(y.shape), dtype=y.dtype, device=y.device) for _ in range(len(op.layers) - 1)]
    # per-image operand maxima of the s stages in an [L, 2, s, B] arena: the merged weight gradients of a layer then run on the
    # kind::f16 kernel without a max pass (ConvNetOp.new_maxima, GradSink.add_wgrad)
    mx_arena = None
    if step_arenas is not None:
        mx_arena = step_arenas[2]
    elif save and op.maxima_ok(y):
        mx_arena = torch.empty((len(op.layers), 2, s, y.shape[0]), dtype=torch.float32, device=y.device)
    for i in range(s):
        if i == 0:
            yi = y
        else:
            if yi_arena is not None:
                yi = yi_arena[i]
            else:
                yi = torch.empty_like(y) if (save or k_bufs is None) else k_bufs[s + (i & 1)]
            _lin(be, yi, [y] + ks, [1.0] + tab.beta[i - 1], dt_dev=dt_dev, dt_from=1, skip=skip)
        if i == 0 and k0 is not None:
            ks.append(k0)
            saved.append(None)
            continue
        kout = k_bufs[i] if k_bufs is not None else None
        r = op.fwd(yi, save=save, out=kout, skip=skip, bufs=[a[i] for a in arenas] if arenas is not None else None,
                   maxima=mx_arena[:, :, i] if mx_arena is not None else None)
        if save:
            ks.append(r[0])
            saved.append(r[1])
        else:
            ks.append(r)
    _lin(be, y_next, [y] + ks, [1.0] + tab.c_sol, dt_dev=dt_dev, dt_from=1, skip=skip)
    return saved if save else None


def rk_step_bwd(be, op: ConvNetOp, tab: Tableau, saved, a_next: torch.Tensor, dt_dev: torch.Tensor, sink: GradSink,
                extra: Sequence[torch.Tensor] = (), out: Optional[torch.Tensor] = None,
                dk_extra: Optional[dict] = None, step_arenas=None) -> torch.Tensor:
    """Adjoint of rk_step_fwd: given a_next = dL/dy_next returns dL/dy (+ sum(extra)), accumulating parameter
    gradients.  dk_i = dt*(c_i a' + sum_{m>i} beta[m-1][i] dy_m);  dy_i = J_i^T dk_i;  dy = a' + sum_i dy_i.
    dk_extra: optional {stage: (tensor, coef)} added into dk_stage (dopri5 dense-output terms)."""
    s = tab.s
    dys: List[Optional[torch.Tensor]] = [None] * s
    L = len(op.layers)
    shape = (s,) + tuple(a_next.shape)
    g_arenas = None
    if step_arenas is not None:                       # (dk [s, B, ...], [per layer [s, B, ...]]) views of rollout-wide arenas
        dk_arena, g_arenas = step_arenas
    else:
        dk_arena = torch.empty(shape, dtype=a_next.dtype, device=a_next.device)
        if L > 1 and all(c.in_channels == a_next.shape[3] for c, _ in op.layers[1:]):
            g_arenas = [torch.empty(shape, dtype=a_next.dtype, device=a_next.device) for _ in range(L - 1)]
    was_deferred, sink.deferred = sink.deferred, True
    for i in range(s - 1, -1, -1):
        ins, coefs = [a_next], [tab.c_sol[i]]
        for m in range(i + 1, s):
            b = tab.beta[m - 1][i]
            if b != 0.0 and dys[m] is not None:
                ins.append(dys[m])
                coefs.append(b)
        if all(c == 0.0 for c in coefs):
            continue                                   # stage does not influence y_next
        dk = dk_arena[i]
        _lin(be, dk, ins, coefs, dt_dev=dt_dev, dt_from=0)
        dys[i] = op.bwd(saved[i], dk, sink, gbufs=[g[i] for g in g_arenas] if g_arenas is not None else None)
    sink.deferred = was_deferred
    if not was_deferred:
        sink.flush()
    terms = [a_next] + [d for d in dys if d is not None] + list(extra)
    res = out if out is not None else torch.empty_like(a_next)
    while len(terms) > 8:                              # odex_lincomb takes <= 8 operands
        part = torch.empty_like(a_next)
        be.lincomb(part, terms[:8], [1.0] * 8)
        terms = [part] + terms[8:]
    be.lincomb(res, terms, [1.0] * len(terms))
    return res


class RolloutArenas:
    """Stage-major arenas [s, n_steps, B, ...] for a fixed-grid rollout under training: stage inputs, saved activations and (in
    the backward) the gradient planes of ALL steps, with their per-image maxima [L, 2, s, n_steps, B] beside them.  A layer's
    weight-gradient requests over all stages AND steps are then consecutive slices of one buffer and run as one launch
    (GradSink._slice_runs): for rk4 on a 10-point grid 6 launches of 1152 - 4608 images instead of 54 of 128 - 512, which
    amortises the kernel's fixed part (pipeline fill + 5 taps of TMEM -> red.global.add epilogue: ~11 us of a 38 us launch).
    Costs nothing extra in memory: the deferred requests keep these planes alive until the flush anyway.  Used when the planes
    of the rollout fit `ODEX_ARENA_GB` (default 16) gigabytes, else the per-step arenas of rk_step_fwd / rk_step_bwd."""

    def __init__(self, op: ConvNetOp, tab: Tableau, y: torch.Tensor, n_steps: int):
        self.ok = False
        L, s = len(op.layers), tab.s
        if s < 2 or L < 2 or n_steps < 2 or y.shape[0] == 0:
            return
        if not all(c.out_channels == y.shape[3] and c.in_channels == y.shape[3] for c, _ in op.layers):
            return
        plane = y.numel() * 4
        import os
        if 2 * L * s * n_steps * plane > float(os.environ.get("ODEX_ARENA_GB", "16")) * 2 ** 30:
            return
        self.ok, self.s, self.n, self.L = True, s, n_steps, L
        self.shape = (s, n_steps) + tuple(y.shape)
        kw = dict(dtype=torch.float32, device=y.device)
        self.yi = torch.empty(self.shape, **kw)
        self.act = [torch.empty(self.shape, **kw) for _ in range(L - 1)]
        self.mx = torch.empty((L, 2, s, n_steps, y.shape[0]), **kw) if op.maxima_ok(y) else None
        self.dk = None
        self.g = None

    def fwd(self, n: int):
        return self.yi[:, n], [a[:, n] for a in self.act], (self.mx[:, :, :, n] if self.mx is not None else None)

    def bwd(self, n: int):
        if self.dk is None:
            kw = dict(dtype=torch.float32, device=self.yi.device)
            self.dk = torch.empty(self.shape, **kw)
            self.g = [torch.empty(self.shape, **kw) for _ in range(self.L - 1)]
        return self.dk[:, n], [g[:, n] for g in self.g]


class _FixedGridFn(torch.autograd.Function):
    """sol[0] = y0; sol[n+1] = RK step from sol[n] with dt = t[n+1]-t[n] (the integration grid IS t)."""

    @staticmethod
    def forward(ctx, op: ConvNetOp, tab: Tableau, y0: torch.Tensor, dts: torch.Tensor, *params):
        be = K.backend()
        be.require(y0)
        y = to_nhwc(y0.detach().float())
        T = dts.numel() + 1
        save = any(ctx.needs_input_grad)
        sol = torch.empty((T,) + tuple(y.shape), dtype=torch.float32, device=y.device)
        be.lincomb(sol[0], [y if y.is_contiguous() else y.contiguous()], [1.0])
        k_bufs = None if save else [torch.empty_like(sol[0]) for _ in range(tab.s + 2)]
        saved = []
        ra = RolloutArenas(op, tab, sol[0], T - 1) if save else None
        if ra is not None and not ra.ok:
            ra = None
        for n in range(T - 1):
            saved.append(rk_step_fwd(be, op, tab, sol[n], sol[n + 1], dts[n:n + 1], save, k_bufs=k_bufs,
                                     step_arenas=ra.fwd(n) if ra is not None else None))
        ctx.op, ctx.tab, ctx.dts, ctx.saved, ctx.params, ctx.ra = op, tab, dts, saved, params, ra
        return from_nhwc(sol)

    @staticmethod
    def backward(ctx, dsol):
        be = K.backend()
        op, tab, dts, saved = ctx.op, ctx.tab, ctx.dts, ctx.saved
        if saved is None:
            raise RuntimeError("DiffEqSolver: this rollout was already back-propagated through once; its saved stage "
                               "activations are freed during that pass (no retain_graph) -- run the forward again")
        d = to_nhwc(dsol.float())
        if not d.is_contiguous():
            d = d.contiguous()
        T = d.shape[0]
        sink = GradSink()
        ra = ctx.ra
        sink.deferred = ra is not None                        # rollout-wide arenas: one flush at the end, merged over all steps
        a = d[T - 1]
        for n in range(T - 2, -1, -1):
            if tab.s == 1:
                dk = torch.empty_like(a)
                be.lincomb(dk, [a], [1.0], dtmask=1, dt_dev=dts[n:n + 1])
                tmp = torch.empty_like(a)
                be.lincomb(tmp, [a, d[n]], [1.0, 1.0])
                a = op.bwd(saved[n][0], dk, sink, res=tmp)
            else:
                a = rk_step_bwd(be, op, tab, saved[n], a, dts[n:n + 1], sink, extra=[d[n]],
                                step_arenas=ra.bwd(n) if ra is not None else None)
            saved[n] = None                                   # free this step's activations
        ctx.saved = ctx.ra = None
        dy0 = from_nhwc(a) if ctx.needs_input_grad[2] else None
        return (None, None, dy0, None) + sink.result(ctx.params)


def check_time_grid(t: torch.Tensor) -> None:
    """odeint._check_inputs: t strictly monotone.  Checked on the host when that costs nothing (t still on the CPU, as
    helpers/utils.py:118-120 creates it) or when no CUDA graph is being captured; a decreasing grid is fine for the
    fixed-grid solvers (torchdiffeq flips the sign of t and f: dt * f is the same number)."""
    if t.numel() < 3 or (t.is_cuda and torch.cuda.is_current_stream_capturing()):
        return
    d = t.detach()[1:] - t.detach()[:-1]
    if not (bool((d > 0).all()) or bool((d < 0).all())):
        raise AssertionError("t must be strictly increasing or decreasing")


def odeint_fixed(op: ConvNetOp, y0: torch.Tensor, t: torch.Tensor, method: str) -> torch.Tensor:
    """euler / rk4 on the output grid.  y0: logical [B,C,H,W]; returns logical [T,B,C,H,W] (channels-last memory)."""
    tab = FIXED[method]
    if not t.is_cuda:
        check_time_grid(t)
    td = t.detach().to(device=y0.device, dtype=torch.float64)
    dts = (td[1:] - td[:-1]).contiguous()
    return _FixedGridFn.apply(op, tab, live_input(y0), dts, *live(op.params()))


# ------------------------------------------------------------------------------------------------- dopri5
DOPRI5 = Tableau("dopri5", DP_BETA, [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0.0])
_ST_F64 = _lib.STATE_HEAD_BYTES // 8
_I_DT_TRY = _lib.OFF_DT_TRY // 8
_I_H0 = _lib.OFF_H0 // 8
_I_RED = _lib.OFF_RED // 8
_I_DONE = _lib.OFF_DONE // 4


class DopriStats:
    """What the device-resident controller reported for one forward (read back once, after the last chunk)."""

    def __init__(self, st: _lib.DopriState, attempts, syncs: int, t_host=None):
        self.n_accept, self.n_reject = st.n_accept, st.n_reject
        self.n_feval, self.n_attempt = st.n_feval, st.n_attempt
        self.status, self.host_syncs = st.status, syncs
        self.t_final, self.dt_final = st.t1, st.dt
        self.attempts = attempts            # [(t0, dt, err_ratio, accepted)] (the first log_cap of them)
        self.t_host = t_host                # the output times, from the same read

    def __repr__(self):
        return "DopriStats(accept=%d reject=%d feval=%d syncs=%d status=%d)" % (
            self.n_accept, self.n_reject, self.n_feval, self.host_syncs, self.status)


def _read_back(ws: "Dopri5Workspace", t_dev: torch.Tensor, issued: int):
    """THE device->host read of a forward (one per speculative chunk): the 256-byte controller header, the output times
    and the log rows of the attempts issued so far, gathered on the device into one buffer and copied once."""
    n_log = max(0, min(issued, ws.log.shape[0]))
    parts = [ws.st[:_lib.STATE_HEAD_BYTES], t_dev.view(torch.uint8)]
    if n_log:
        parts.append(ws.log[:n_log].view(torch.uint8).reshape(-1))
    host = torch.cat(parts).cpu()
    raw = host.numpy().tobytes()
    st = _lib.DopriState.from_buffer_copy(raw[:ctypes.sizeof(_lib.DopriState)])
    o = _lib.STATE_HEAD_BYTES
    T = t_dev.numel()
    t_host = host[o:o + 8 * T].view(torch.float64).tolist()
    rows = host[o + 8 * T:].view(torch.float64).reshape(-1, 4).tolist() if n_log else []
    return st, t_host, [tuple(r) for r in rows[:min(st.n_attempt, n_log)]]


class Dopri5Workspace:
    """Device buffers of one adaptive integration (state vector y, 7 stage planes, controller state, log)."""

    def __init__(self, like: torch.Tensor, n_out: int, log_cap: int, save_cap: int, be):
        dev = like.device
        self.y = torch.empty_like(like)
        self.y1 = torch.empty_like(like)
        self.ystage = [torch.empty_like(like) for _ in range(2)]
        self.k = [torch.empty_like(like) for _ in range(7)]
        self.st = torch.zeros(be.dopri_state_bytes(), dtype=torch.uint8, device=dev)
        self.st64 = self.st[:_lib.STATE_HEAD_BYTES].view(torch.float64)
        self.st32 = self.st[:_lib.STATE_HEAD_BYTES].view(torch.int32)
        self.dt_try = self.st64[_I_DT_TRY:_I_DT_TRY + 1]
        self.h0 = self.st64[_I_H0:_I_H0 + 1]
        self.red = self.st64[_I_RED:_I_RED + 4]
        self.done = self.st32[_I_DONE:_I_DONE + 1]
        self.log = torch.zeros((log_cap, 4), dtype=torch.float64, device=dev)
        self.ysave = self.fsave = None
        self.save_cap = save_cap
        if save_cap > 0:
            self.ysave = torch.empty((save_cap,) + tuple(like.shape), dtype=torch.float32, device=dev)
            self.fsave = torch.empty((save_cap,) + tuple(like.shape), dtype=torch.float32, device=dev)


def dopri5_forward(op: ConvNetOp, y: torch.Tensor, t_dev: torch.Tensor, rtol: float, atol: float,
                   reduce_fn=None, chunk_hint: int = 8, max_attempts: int = 4096, save_cap: int = 0,
                   log_cap: Optional[int] = None):
    """Adaptive Dormand-Prince 5(4) (torchdiffeq rk_common / dopri5 / interp) over dense NHWC
    state y.  Returns (sol [T, ...], stats, workspace).  Attempts are enqueued speculatively in chunks; every
    kernel of an attempt is a no-op once the device `done` flag is set, so a chunk that overshoots only costs
    empty launches.  reduce_fn(red): optional in-stream all-reduce(SUM) of the 4 fp64 reduction slots, so that
    a batch sharded over ranks takes the step sequence of the global batch."""
    be = K.backend()
    T = t_dev.numel()
    if log_cap is None:
        log_cap = max(1, min(int(max_attempts), 1 << 16))
    ws = Dopri5Workspace(y, T, log_cap, save_cap, be)
    sol = torch.empty((T,) + tuple(y.shape), dtype=torch.float32, device=y.device)
    be.lincomb(sol[0], [y], [1.0])
    be.lincomb(ws.y, [y], [1.0])
    be.dopri_init(ws.st, t_dev, T, rtol, atol, max_attempts, save_cap)
    syncs = 0
    issued = 0
    fused = reduce_fn is None          # single GPU: the norm kernels' last block runs the controller (no 1-thread launches)
    if T > 1:
        # ---- initial step size (misc._select_initial_step, order 4): two f-evals and three RMS norms
        op.fwd(ws.y, out=ws.k[0])
        be.dopri_initnorm(ws.st, 0, ws.y, ws.k[0], None, fuse_ctrl=fused)
        if not fused:
            reduce_fn(ws.red)
            be.dopri_initctrl(ws.st, 0)
        be.lincomb(ws.y1, [ws.y, ws.k[0]], [1.0, 1.0], dtmask=2, dt_dev=ws.h0)
        op.fwd(ws.y1, out=ws.k[1])
        be.dopri_initnorm(ws.st, 1, ws.y, ws.k[0], ws.k[1], fuse_ctrl=fused)
        if not fused:
            reduce_fn(ws.red)
            be.dopri_initctrl(ws.st, 1)
        # ---- attempts
        chunk = max(1, int(chunk_hint))
        while True:
            for _ in range(chunk):
                _dopri_attempt(be, op, ws, t_dev, sol, reduce_fn, issued)
                issued += 1
            st, t_host, attempts = _read_back(ws, t_dev, issued)
            syncs += 1
            if st.done or issued >= max_attempts:
                break
            chunk = max(2, min(64, st.n_attempt // 2))
    else:
        st, t_host, attempts = _read_back(ws, t_dev, 0)
        syncs += 1
    stats = DopriStats(st, attempts, syncs, t_host)
    # torchdiffeq's asserts (odeint._check_inputs, rk_common._adaptive_step), raised once per forward from the host copy of t
    # and the device-side sticky flags
    if any(not (b > a) for a, b in zip(t_host[:-1], t_host[1:])):
        raise AssertionError("t must be strictly increasing (dopri5 on a decreasing or unsorted time grid is not supported)")
    if st.status & 1:
        raise AssertionError("underflow in dt {}".format(st.dt))
    if st.status & 2:
        raise AssertionError("non-finite values in state `y`")
    if st.status & 8:
        raise RuntimeError("dopri5: more than %d accepted steps -- the backward checkpoint capacity; the integration was "
                           "stopped there.  Raise DiffEqSolver.save_budget_bytes" % save_cap)
    if st.status & 4 or not st.done:
        raise AssertionError("max_num_steps exceeded ({} attempts)".format(st.n_attempt))
    return sol, stats, ws


def _dopri_attempt(be, op: ConvNetOp, ws: Dopri5Workspace, t_dev, sol, reduce_fn, seq: int):
    """One speculative attempt: 6 x (stage combination + f-eval), then K4 (error estimate + RMS norm + accept/reject +
    step-size controller, one kernel) and K5 (dense output + FSAL advance): 14 launches, none of them single-threaded.
    With a sharded batch the 4 partial sums are all-reduced between the norm and the decision (one more small launch)."""
    skip = ws.done
    for s in range(6):
        dst = ws.y1 if s == 5 else ws.ystage[s & 1]
        _lin(be, dst, [ws.y] + ws.k[:s + 1], [1.0] + DP_BETA[s], dt_dev=ws.dt_try, dt_from=1, skip=skip)
        op.fwd(dst, out=ws.k[s + 1], skip=skip)
    if reduce_fn is None:
        be.dopri_errctrl(ws.st, t_dev, ws.y, ws.y1, ws.k, ws.log, ws.log.shape[0], seq)
    else:
        be.dopri_errnorm(ws.st, ws.y, ws.y1, ws.k)
        reduce_fn(ws.red)
        be.dopri_ctrl(ws.st, t_dev, ws.log, ws.log.shape[0], seq)
    be.dopri_advance(ws.st, t_dev, ws.y, ws.y1, ws.k, sol, ws.ysave, ws.fsave, ws.save_cap, seq)


def _interp_polys(x: float):
    """Quartic dense output (interp._interp_fit/_interp_evaluate) as weights of (y0, y1, y_mid, dt*f0, dt*f1)."""
    x2, x3, x4 = x * x, x * x * x, x * x * x * x
    return (1 - 11 * x2 + 18 * x3 - 8 * x4, -5 * x2 + 14 * x3 - 8 * x4, 16 * x2 - 32 * x3 + 16 * x4,
            x - 4 * x2 + 5 * x3 - 2 * x4, x2 - 3 * x3 + 2 * x4)


def _sum_into(be, out, ins, coefs):
    """out = sum coefs*ins for any number of operands (chunks of 8)."""
    ins, coefs = list(ins), list(coefs)
    first = True
    while ins:
        if first:
            a, c = ins[:8], coefs[:8]
            ins, coefs = ins[8:], coefs[8:]
        else:
            a, c = [out] + ins[:7], [1.0] + coefs[:7]
            ins, coefs = ins[7:], coefs[7:]
        be.lincomb(out, a, c)
        first = False
    return out


class _Dopri5Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, op: ConvNetOp, solver, y0: torch.Tensor, t_dev: torch.Tensor, rtol, atol, *params):
        be = K.backend()
        be.require(y0)
        y = to_nhwc(y0.detach().float())
        if not y.is_contiguous():
            y = y.contiguous()
        save = any(ctx.needs_input_grad)
        save_cap = 0
        if save:
            budget = int(getattr(solver, "save_budget_bytes", 8 << 30))
            save_cap = max(4, min(512, budget // max(1, 2 * y.numel() * 4)))
        sol, stats, ws = dopri5_forward(op, y, t_dev, rtol, atol, reduce_fn=getattr(solver, "reduce_fn", None),
                                        chunk_hint=getattr(solver, "_chunk_hint", 8),
                                        max_attempts=getattr(solver, "max_attempts", 4096), save_cap=save_cap)
        solver._chunk_hint = stats.n_attempt + 1        # next forward: one chunk, one host sync
        solver.last_stats = stats
        if save:
            if stats.n_attempt > len(stats.attempts):
                raise RuntimeError("dopri5: %d attempts exceed the step log (%d rows) the backward replays; lower "
                                   "DiffEqSolver.max_attempts" % (stats.n_attempt, len(stats.attempts)))
            ctx.op, ctx.ws, ctx.stats, ctx.params = op, ws, stats, params
            ctx.t_host = stats.t_host               # from the forward's single read-back
        return from_nhwc(sol)

    @staticmethod
    def backward(ctx, dsol):
        be = K.backend()
        op, ws, stats = ctx.op, ctx.ws, ctx.stats
        if ws is None:
            raise RuntimeError("DiffEqSolver(dopri5): this rollout was already back-propagated through once; its step "
                               "checkpoints are freed during that pass (no retain_graph) -- run the forward again")
        d = to_nhwc(dsol.float())
        if not d.is_contiguous():
            d = d.contiguous()
        t = ctx.t_host
        T = len(t)
        sink = GradSink()
        like = d[0]
        new = lambda: torch.empty_like(like)   # noqa: E731
        # accepted attempts and the outputs each one produced (same fp64 arithmetic as dopri_ctrl_kernel)
        acc = [(i, a[0], a[1]) for i, a in enumerate(stats.attempts) if a[3] != 0.0]
        steps, j = [], 1
        for (ai, t0, dt) in acc:
            t1 = t0 + dt
            js = []
            while j < T and not (t[j] > t1):
                js.append(j)
                j += 1
            steps.append((ai, t0, t1, dt, js))
        A_y = None      # adjoint of the state at the END of the step being processed
        A_f = None      # adjoint of f at the end of the step (FSAL: the next step's k0)
        for m in range(len(steps) - 1, -1, -1):
            ai, t0, t1, dt, js = steps[m]
            dt_dev = ws.log[ai, 1:2]
            y, k0 = ws.ysave[m], ws.fsave[m]
            # ---- recompute the six stages of this accepted step, keeping the conv activations -- in [6, B, ...] arenas (stage
            # inputs, activations, gradient planes, per-image maxima), so that a layer's six weight gradients of the step run as
            # one launch at the step's end (GradSink.add_wgrad)
            L_ = len(op.layers)
            ar = None
            if L_ > 1 and all(c.out_channels == like.shape[3] and c.in_channels == like.shape[3] for c, _ in op.layers):
                mk = lambda: torch.empty((6,) + tuple(like.shape), dtype=like.dtype, device=like.device)   # noqa: E731
                ar = {"yi": mk(), "act": [mk() for _ in range(L_ - 1)], "dk": mk(), "g": [mk() for _ in range(L_ - 1)],
                      "mx": torch.empty((L_, 2, 6, like.shape[0]), dtype=torch.float32, device=like.device) if op.maxima_ok(like) else None}
                sink.deferred = True
            ks, saved = [k0], [None]
            for s in range(6):
                yi = ar["yi"][s] if ar is not None else new()
                _lin(be, yi, [y] + ks, [1.0] + DP_BETA[s], dt_dev=dt_dev, dt_from=1)
                if ar is not None:
                    ki, sv = op.fwd(yi, save=True, bufs=[a_[s] for a_ in ar["act"]],
                                    maxima=ar["mx"][:, :, s] if ar["mx"] is not None else None)
                else:
                    ki, sv = op.fwd(yi, save=True, maxima=op.new_maxima(yi))
                ks.append(ki)
                saved.append(sv)
            # ---- dense-output adjoints: sol_j = P(x_j) . (y0, y1, y_mid, dt f0, dt f1)
            E = {}
            if js:
                G = [d[jj] for jj in js]
                P = [_interp_polys(float(torch.tensor((t[jj] - t0) / (t1 - t0), dtype=torch.float64).float()))
                     for jj in js]
                E["y0"] = _sum_into(be, new(), G, [p[0] + p[2] for p in P])
                E["y1"] = _sum_into(be, new(), G, [p[1] for p in P])
                dtf = float(torch.tensor(dt, dtype=torch.float64).float())
                for i in (0, 2, 3, 4, 5, 6):
                    cf = [dtf * p[2] * DP_C_MID[i] + (dtf * p[3] if i == 0 else 0.0) + (dtf * p[4] if i == 6 else 0.0)
                          for p in P]
                    E[i] = _sum_into(be, new(), G, cf)
            # ---- stage 6: k6 = f(y1) feeds the next step (FSAL) and the interpolant
            terms = [x for x in (A_f, E.get(6)) if x is not None]
            Ay6_terms = [x for x in (A_y, E.get("y1")) if x is not None]
            dys: List[Optional[torch.Tensor]] = [None] * 7
            if terms:
                dk6 = terms[0] if len(terms) == 1 else _sum_into(be, new(), terms, [1.0] * len(terms))
                dys[6] = op.bwd(saved[6], dk6, sink, gbufs=[g_[5] for g_ in ar["g"]] if ar is not None else None)
                Ay6_terms.append(dys[6])
            if not Ay6_terms:
                A_y, A_f = None, None
                sink.flush()
                sink.deferred = False
                continue
            Ay6 = Ay6_terms[0] if len(Ay6_terms) == 1 else _sum_into(be, new(), Ay6_terms, [1.0] * len(Ay6_terms))
            # ---- stages 5..1, then k0 (an input of the step: its adjoint goes to the previous step's k6)
            dk0 = None
            for i in range(5, -1, -1):
                ins, coefs = [Ay6], [DP_BETA[5][i]]
                for mm in range(i + 1, 6):
                    if dys[mm] is not None and DP_BETA[mm - 1][i] != 0.0:
                        ins.append(dys[mm])
                        coefs.append(DP_BETA[mm - 1][i])
                keep = [(a_, c_) for a_, c_ in zip(ins, coefs) if c_ != 0.0]
                extra = E.get(i)
                if not keep and extra is None:
                    continue
                dk = ar["dk"][i - 1] if (ar is not None and i > 0) else new()
                if keep:
                    n_dt = len(keep)
                    ops_ = [a_ for a_, _ in keep] + ([extra] if extra is not None else [])
                    cfs_ = [c_ for _, c_ in keep] + ([1.0] if extra is not None else [])
                    be.lincomb(dk, ops_, cfs_, dtmask=(1 << n_dt) - 1, dt_dev=dt_dev)
                else:
                    be.lincomb(dk, [extra], [1.0])
                if i == 0:
                    dk0 = dk
                else:
                    dys[i] = op.bwd(saved[i], dk, sink, gbufs=[g_[i - 1] for g_ in ar["g"]] if ar is not None else None)
            sink.flush()                                      # the step's weight gradients, merged over its stages
            sink.deferred = False
            terms = [Ay6] + [dys[i] for i in range(1, 6) if dys[i] is not None]
            if "y0" in E:
                terms.append(E["y0"])
            A_y = _sum_into(be, new(), terms, [1.0] * len(terms))
            A_f = dk0
        # ---- the very first f-eval k0 = f(y0), and sol[0] = y0
        y_first = ws.ysave[0] if steps else None
        terms = [d[0]] + ([A_y] if A_y is not None else [])
        if A_f is not None and y_first is not None:
            _, sv = op.fwd(y_first, save=True)
            terms.append(op.bwd(sv, A_f, sink))
        dy0 = _sum_into(be, new(), terms, [1.0] * len(terms))
        ctx.ws = None
        return (None, None, from_nhwc(dy0) if ctx.needs_input_grad[2] else None, None, None, None) + sink.result(ctx.params)


def odeint_dopri5(op: ConvNetOp, solver, y0: torch.Tensor, t: torch.Tensor, rtol: float, atol: float) -> torch.Tensor:
    td = t.detach().to(device=y0.device, dtype=torch.float64).contiguous()
    return _Dopri5Fn.apply(op, solver, live_input(y0), td, float(rtol), float(atol), *live(op.params()))
