"""TEST INFRASTRUCTURE ONLY -- CPU restatement of torchdiffeq==0.2.1 `odeint`.

PARITY UNPINNED: the reference (`/root/reference/modules/DiffEqSolver.py:9`) does
`from torchdiffeq import odeint`, pinned `torchdiffeq==0.2.1`
(`/root/reference/requirements.txt:102`).  The package is third-party, not vendored in the
reference tree, not installed in this image and not downloadable (no network).  This file
restates the published algorithm of that release (odeint.py, misc.py, solvers.py,
fixed_grid.py, rk_common.py, dopri5.py, interp.py) as specified in SURVEY.md Appendix A.
The reference holds no tests / golden vectors for the path, so this oracle is pinned only by
(i) its self-tests in tests/test_oracle_odeint.py -- closed-form solutions, measured convergence
orders 1 / 4 / 5, and an INDEPENDENT implementation of the same published method: scipy 1.x
`integrate._ivp.rk.RK45` (same Dormand-Prince tableau and Shampine dense output; single steps,
the error estimate up to the known 2/3 factor between the two embedded pairs, the dense output
and `select_initial_step` agree to fp64 rounding; whole solves to O(tol)) -- and (ii) the
reference's call sites (`modules/DiffEqSolver.py:37,45-46`).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module.  The product (ode_rl_b200/) never does.

Semantics kept on purpose (they differ from a "textbook" solver):
  * rk4 is the 3/8-rule variant (`rk4_alt_step_func`), not classical RK4.
  * fixed-grid solvers integrate on the output grid `t` itself (no step_size option passed).
  * dopri5: fp64 time / dt, fp32 state; batch-wide RMS norm; steps are NOT clipped to the output
    times (overshoot + quartic interpolation); accepted steps never shrink (dfactor=1 if err<1).
  * `k` stage operands are summed j=0..i ascending in fp32.
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional

import torch

# --- Dormand-Prince 5(4) tableau (torchdiffeq dopri5.py) -------------------------------------
DP_ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
DP_BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
DP_C_SOL = [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0.0]
DP_C_ERROR = [
    35 / 384 - 1951 / 21600,
    0.0,
    500 / 1113 - 22642 / 50085,
    125 / 192 - 451 / 720,
    -2187 / 6784 - -12231 / 42400,
    11 / 84 - 649 / 6300,
    -1.0 / 60.0,
]
DP_C_MID = [
    6025192743 / 30085553152 / 2,
    0.0,
    51252292925 / 65400821598 / 2,
    -2691868925 / 45128329728 / 2,
    187940372067 / 1594534317056 / 2,
    -1776094331 / 19743644256 / 2,
    11237099 / 235043384 / 2,
]


def _rms_norm(x: torch.Tensor) -> torch.Tensor:
    # misc.py `_rms_norm`: over the WHOLE tensor (batch-shared step size)
    return x.pow(2).mean().sqrt()


class StepLog:
    """Per-attempt record for dopri5: (t0, dt, err_ratio, accepted)."""

    def __init__(self):
        self.attempts: List[tuple] = []
        self.n_feval = 0
        self.dt0: Optional[float] = None

    @property
    def n_accept(self):
        return sum(1 for a in self.attempts if a[3])

    @property
    def n_reject(self):
        return sum(1 for a in self.attempts if not a[3])


# --- fixed grid ------------------------------------------------------------------
def _euler_step(func, t0, dt, y):
    return dt * func(t0, y)


def _rk4_alt_step(func, t0, dt, y):
    k1 = func(t0, y)
    k2 = func(t0 + dt / 3, y + dt * k1 * (1.0 / 3.0))
    k3 = func(t0 + dt * 2 / 3, y + dt * (k2 - k1 * (1.0 / 3.0)))
    k4 = func(t0 + dt, y + dt * (k1 - k2 + k3))
    return (k1 + 3 * (k2 + k3) + k4) * dt * 0.125


def _fixed_grid(func, y0, t, step):
    sol = [y0]
    y = y0
    for t0, t1 in zip(t[:-1], t[1:]):
        dt = t1 - t0                      # 0-dim fp64; fp32 tensor * 0-dim fp64 stays fp32
        dy = step(func, t0, dt, y)
        y = y + dy
        sol.append(y)
    return torch.stack(sol, 0)


# --- dopri5 ----------------------------------------------------------------
def _select_initial_step(func, t0, y0, order, rtol, atol, f0, log):
    scale = atol + y0.abs() * rtol
    d0 = _rms_norm(y0 / scale)
    d1 = _rms_norm(f0 / scale)
    if d0 < 1e-5 or d1 < 1e-5:
        h0 = torch.tensor(1e-6, dtype=y0.dtype)
    else:
        h0 = 0.01 * d0 / d1
    y1 = y0 + h0 * f0
    f1 = func(t0 + h0, y1)
    log.n_feval += 1
    d2 = _rms_norm((f1 - f0) / scale) / h0
    if d1 <= 1e-15 and d2 <= 1e-15:
        h1 = torch.max(torch.tensor(1e-6, dtype=y0.dtype), h0 * 1e-3)
    else:
        h1 = (0.01 / max(d1, d2)) ** (1.0 / float(order + 1))
    return torch.min(100 * h0, h1).to(torch.float64)


def _optimal_step_size(last_step, error_ratio, safety=0.9, ifactor=10.0, dfactor=0.2, order=5):
    # misc.py `_optimal_step_size` (no_grad, time dtype)
    error_ratio = float(error_ratio)
    if error_ratio == 0:
        return last_step * ifactor
    if error_ratio < 1:
        dfactor = 1.0
    exponent = 1.0 / order
    factor = min(ifactor, max(safety / error_ratio ** exponent, dfactor))
    return last_step * factor


def _interp_fit(y0, y1, y_mid, f0, f1, dt):
    a = 2 * dt * (f1 - f0) - 8 * (y1 + y0) + 16 * y_mid
    b = dt * (5 * f0 - 3 * f1) + 18 * y0 + 14 * y1 - 32 * y_mid
    c = dt * (f1 - 4 * f0) - 11 * y0 - 5 * y1 + 16 * y_mid
    d = dt * f0
    e = y0
    return [e, d, c, b, a]


def _interp_evaluate(coeffs, t0, t1, t):
    x = ((t - t0) / (t1 - t0)).to(coeffs[0].dtype)
    total = coeffs[0] + x * coeffs[1]
    x_power = x
    for c in coeffs[2:]:
        x_power = x_power * x
        total = total + x_power * c
    return total


def _runge_kutta_step(func, y, f, t0, dt, log=None):
    """rk_common._runge_kutta_step with the Dormand-Prince tableau: (y1, f1, y1_error, k[0..6]).  t0 / dt fp64 0-dim
    tensors (or floats); the stage sums run j = 0..i ascending in the state dtype, as `k[..., :i+1].matmul(beta_i * dt)`."""
    dtype = y.dtype
    dt32 = torch.as_tensor(dt, dtype=torch.float64).to(dtype)
    k = [f]
    yi = y
    for s in range(6):
        acc = None
        for j in range(s + 1):
            term = k[j] * (dt32 * torch.tensor(DP_BETA[s][j], dtype=dtype))
            acc = term if acc is None else acc + term
        yi = y + acc
        k.append(func(t0 + DP_ALPHA[s] * dt, yi))
        if log is not None:
            log.n_feval += 1
    y1, f1 = yi, k[6]                       # FSAL: c_sol[:-1] == beta[5] and c_sol[6] == 0
    err = None
    for j in range(7):
        term = k[j] * (dt32 * torch.tensor(DP_C_ERROR[j], dtype=dtype))
        err = term if err is None else err + term
    return y1, f1, err, k


def _dopri5_interp_fit(y, y1, k, dt):
    """dopri5.Dopri5Solver._interp_fit: y_mid from c_mid, then the quartic through (y0, f0, y_mid, y1, f1)."""
    dtype = y.dtype
    dt32 = torch.as_tensor(dt, dtype=torch.float64).to(dtype)
    mid = None
    for j in range(7):
        term = k[j] * (dt32 * torch.tensor(DP_C_MID[j], dtype=dtype))
        mid = term if mid is None else mid + term
    y_mid = y + mid
    return _interp_fit(y, y1, y_mid, k[0], k[6], dt32)


def _dopri5(func, y0, t, rtol, atol, log, detach_dt0=True, max_num_steps=2 ** 31 - 1):  # noqa
    t = t.to(torch.float64)
    f0 = func(t[0], y0)
    log.n_feval += 1
    with torch.no_grad() if detach_dt0 else _nullctx():
        dt = _select_initial_step(func, t[0], y0, 4, rtol, atol, f0, log)
    if detach_dt0:
        dt = dt.detach()
    log.dt0 = float(dt)
    y, f, t0_prev, t1 = y0, f0, t[0], t[0]
    coeffs = [y0] * 5
    sol = [y0]
    n_steps = 0
    for i in range(1, len(t)):
        while t[i] > t1:
            assert n_steps < max_num_steps, "max_num_steps exceeded"
            assert t1 + dt > t1, "underflow in dt {}".format(float(dt))
            assert torch.isfinite(y).all(), "non-finite values in state `y`: {}".format(y)
            t0_a = t1
            y1, f1, err, k = _runge_kutta_step(func, y, f, t0_a, dt, log)
            tol = atol + rtol * torch.max(y.abs(), y1.abs())
            ratio = _rms_norm(err / tol)
            accept = bool(ratio <= 1)
            log.attempts.append((float(t0_a), float(dt), float(ratio.detach()), accept))
            if accept:
                coeffs = _dopri5_interp_fit(y, y1, k, dt)
                t0_prev, t1 = t0_a, t0_a + dt
                y, f = y1, f1
            dt = _optimal_step_size(dt, ratio.detach().to(torch.float64))
            n_steps += 1
        sol.append(_interp_evaluate(coeffs, t0_prev, t1, t[i]))
    return torch.stack(sol, 0)


class _nullctx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# --- public entry ----------------------------------------------------------------
LAST_LOG: Optional[StepLog] = None


def odeint(func: Callable, y0: torch.Tensor, t: torch.Tensor, *, rtol=1e-7, atol=1e-9,
           method: Optional[str] = None, options=None, log: Optional[StepLog] = None,
           detach_dt0: bool = True):
    """torchdiffeq 0.2.1 `odeint` for method in {euler, rk4, dopri5}; returns [len(t), *y0.shape]."""
    global LAST_LOG
    assert t.ndim == 1, "t must be one dimensional"
    assert bool((t[1:] > t[:-1]).all()) or len(t) == 1, "t must be strictly increasing (path only uses increasing t)"
    if method is None:
        method = "dopri5"
    if log is None:
        log = StepLog()
    LAST_LOG = log
    if len(t) == 1:
        return y0.unsqueeze(0).clone()
    if method == "euler":
        return _fixed_grid(func, y0, t, _euler_step)
    if method == "rk4":
        return _fixed_grid(func, y0, t, _rk4_alt_step)
    if method == "dopri5":
        return _dopri5(func, y0, t, rtol, atol, log, detach_dt0=detach_dt0)
    raise ValueError("Invalid method \"{}\" (oracle restates euler / rk4 / dopri5 only)".format(method))
