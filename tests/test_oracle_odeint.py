"""Self-tests that PIN oracle/torchdiffeq_021.py -- the CPU restatement of the absent third-party solver the reference
calls (`from torchdiffeq import odeint`, modules/DiffEqSolver.py:9,45-46; torchdiffeq==0.2.1, requirements.txt:102) --
against things that are NOT this repo (VERDICT r1 "next" item 1a):

  * closed-form solutions (y' = lambda*y, harmonic oscillator) and measured convergence orders 1 / 4 / 5;
  * scipy's RK45 (scipy.integrate._ivp.rk / common), an independent implementation of the same published method
    (Dormand-Prince 5(4) tableau, Hairer's initial-step heuristic, Shampine's quartic dense output):
      - tableau identities: beta == A, c_sol == B, c_error == -(2/3)*E (torchdiffeq inherited the tf.contrib.integrate
        embedded pair, whose error estimate is exactly 2/3 of Hairer's), c_mid == the scipy dense output at x = 1/2;
      - one step: y1, f1, the error estimate and the dense output at any x agree to fp64 rounding;
      - select_initial_step: identical value;
      - whole solves: agreement to O(tol) (the two step-size controllers differ, so the step sequences do);
  * structural identities the kernels rely on: FSAL (c_sol[:-1] == beta[5], c_sol[6] == 0), the 3/8-rule RK4, the
    controller's clamp rules, len(t) == 1 -> y0 (the `memory=True` degenerate case).
"""
import math

import numpy as np
import pytest
import torch

from oracle import torchdiffeq_021 as tde

scipy_rk = pytest.importorskip("scipy.integrate._ivp.rk")
scipy_common = pytest.importorskip("scipy.integrate._ivp.common")
from scipy.integrate import solve_ivp  # noqa: E402

F64 = torch.float64


def _osc(t, y):
    """y = (q, p): q' = p, p' = -q."""
    return torch.stack([y[1], -y[0]])


def _nonlin(t, y):
    """A smooth NON-linear autonomous field (the order conditions of a linear problem are a strict subset)."""
    return torch.stack([torch.sin(y[1]) + 0.3 * y[0] * y[1], torch.cos(y[0]) - 0.5 * y[1] ** 2 * 0.2])


def _nonlin_np(t, y):
    return np.array([math.sin(y[1]) + 0.3 * y[0] * y[1], math.cos(y[0]) - 0.5 * y[1] ** 2 * 0.2])


# ------------------------------------------------------------------------------------------------ closed forms
@pytest.mark.parametrize("method,tol", [("euler", 2e-2), ("rk4", 1e-9), ("dopri5", 1e-6)])
def test_exponential_decay_matches_closed_form(method, tol):
    lam = -1.3
    y0 = torch.tensor([1.0, -2.0, 0.5], dtype=F64)
    t = torch.linspace(0.0, 1.0, 101, dtype=F64)
    sol = tde.odeint(lambda t_, y: lam * y, y0, t, rtol=1e-8, atol=1e-10, method=method)
    exact = y0[None] * torch.exp(lam * t)[:, None]
    assert sol.shape == exact.shape and torch.equal(sol[0], y0)
    assert float((sol - exact).abs().max()) < tol


@pytest.mark.parametrize("method,tol", [("rk4", 2e-7), ("dopri5", 1e-6)])
def test_harmonic_oscillator_matches_closed_form(method, tol):
    y0 = torch.tensor([1.0, 0.0], dtype=F64)
    t = torch.linspace(0.0, 2 * math.pi, 201, dtype=F64)
    sol = tde.odeint(_osc, y0, t, rtol=1e-8, atol=1e-10, method=method)
    exact = torch.stack([torch.cos(t), -torch.sin(t)], 1)
    assert float((sol - exact).abs().max()) < tol


# ------------------------------------------------------------------------------------------------ convergence order
def _global_error(step_fn, n):
    y = torch.tensor([0.7, -0.4], dtype=F64)
    h = 1.0 / n
    for i in range(n):
        y = step_fn(y, i * h, h)
    return y


def _order(step_fn, ns):
    ref = solve_ivp(_nonlin_np, (0.0, 1.0), [0.7, -0.4], method="DOP853", rtol=1e-13, atol=1e-14).y[:, -1]
    errs = [float(np.abs(_global_error(step_fn, n).numpy() - ref).max()) for n in ns]
    return [math.log(errs[i] / errs[i + 1]) / math.log(ns[i + 1] / ns[i]) for i in range(len(ns) - 1)], errs


def test_convergence_order_euler_rk4_dopri5():
    """Global error ~ h^p on a non-linear problem: p = 1 (euler), 4 (rk4, 3/8 rule), 5 (dopri5's propagated solution)."""
    euler = lambda y, t0, h: y + tde._euler_step(_nonlin, t0, h, y)          # noqa: E731
    rk4 = lambda y, t0, h: y + tde._rk4_alt_step(_nonlin, t0, h, y)          # noqa: E731

    def dp(y, t0, h):
        return tde._runge_kutta_step(_nonlin, y, _nonlin(t0, y), t0, h)[0]

    p, _ = _order(euler, [64, 128, 256])
    assert all(abs(x - 1.0) < 0.06 for x in p), p
    p, _ = _order(rk4, [8, 16, 32])
    assert all(abs(x - 4.0) < 0.25 for x in p), p
    p, errs = _order(dp, [8, 16, 32])
    assert all(abs(x - 5.0) < 0.4 for x in p), (p, errs)


def test_dopri5_error_estimate_is_fifth_order_in_h():
    """The embedded estimate y1_error = dt * sum c_error[j] k_j is the difference of a 5th- and a 4th-order solution: O(h^5)."""
    y = torch.tensor([0.7, -0.4], dtype=F64)
    f = _nonlin(0.0, y)
    e = [float(tde._runge_kutta_step(_nonlin, y, f, 0.0, h)[2].abs().max()) for h in (0.2, 0.1, 0.05)]
    p = [math.log(e[i] / e[i + 1]) / math.log(2.0) for i in range(2)]
    assert all(abs(x - 5.0) < 0.3 for x in p), (p, e)


def test_fixed_grid_integrates_on_the_output_grid_itself():
    """No sub-stepping: one step per output interval (fixed_grid.FixedGridODESolver with no step_size option)."""
    calls = []

    def f(t_, y):
        calls.append(float(t_))
        return -y

    t = torch.tensor([0.0, 0.5, 0.6, 2.0], dtype=F64)
    y0 = torch.ones(3, dtype=F64)
    sol = tde.odeint(f, y0, t, method="euler")
    assert len(calls) == 3
    y = y0
    for a, b in zip(t[:-1], t[1:]):
        y = y + (b - a) * (-y)
    assert torch.allclose(sol[-1], y, rtol=0, atol=1e-15)
    calls.clear()
    tde.odeint(f, y0, t, method="rk4")
    assert len(calls) == 12


def test_rk4_is_the_three_eighths_rule_not_classical_rk4():
    """rk_common.rk4_alt_step_func: one step on y' = t-free nonlinear f equals the Butcher 3/8 tableau evaluated by hand and
    differs from the classical 1/6-2/6-2/6-1/6 rule."""
    y = torch.tensor([0.7, -0.4], dtype=F64)
    h = 0.3
    k1 = _nonlin(0, y)
    k2 = _nonlin(0, y + h * k1 / 3)
    k3 = _nonlin(0, y + h * (-k1 / 3 + k2))
    k4 = _nonlin(0, y + h * (k1 - k2 + k3))
    want = y + h * (k1 + 3 * k2 + 3 * k3 + k4) / 8
    got = y + tde._rk4_alt_step(_nonlin, 0.0, h, y)
    assert float((got - want).abs().max()) < 1e-15
    c2 = _nonlin(0, y + h * k1 / 2)
    c3 = _nonlin(0, y + h * c2 / 2)
    c4 = _nonlin(0, y + h * c3)
    classical = y + h * (k1 + 2 * c2 + 2 * c3 + c4) / 6
    assert float((got - classical).abs().max()) > 1e-7


# ------------------------------------------------------------------------------------------------ scipy RK45 cross-checks
def test_tableau_equals_scipy_rk45():
    RK45 = scipy_rk.RK45
    A = np.zeros((6, 5))
    for i, row in enumerate(tde.DP_BETA[:5]):
        A[i + 1, :len(row)] = row
    assert np.abs(A - RK45.A).max() < 1e-16
    assert np.abs(np.array(tde.DP_BETA[5]) - RK45.B).max() < 1e-16
    assert np.abs(np.array([0.0] + tde.DP_ALPHA[:5]) - RK45.C).max() < 1e-16
    # FSAL identity the kernels rely on (rk_common._runge_kutta_step skips the c_sol combine when it holds)
    assert tde.DP_C_SOL[:-1] == tde.DP_BETA[5] and tde.DP_C_SOL[6] == 0.0
    # embedded pair: torchdiffeq's c_error is exactly -(2/3) x scipy's E (Hairer's pair)
    assert np.abs(np.array(tde.DP_C_ERROR) + (2.0 / 3.0) * RK45.E).max() < 1e-16
    # the midpoint formula c_mid is Shampine's dense output at x = 1/2
    x = 0.5
    w = x * (RK45.P @ np.array([1.0, x, x * x, x ** 3]))
    assert np.abs(w - np.array(tde.DP_C_MID)).max() < 1e-15
    assert abs(sum(tde.DP_C_SOL) - 1.0) < 1e-15 and abs(sum(tde.DP_C_ERROR)) < 1e-15


@pytest.mark.parametrize("h", [0.37, 0.05])
def test_one_step_equals_scipy_rk_step(h):
    """y1, f1 and the error estimate of one attempt against scipy.integrate._ivp.rk.rk_step (fp64 on both sides)."""
    RK45 = scipy_rk.RK45
    y = np.array([0.7, -0.4])
    f = _nonlin_np(0.0, y)
    K = np.empty((7, 2))
    y_new, f_new = scipy_rk.rk_step(_nonlin_np, 0.0, y, f, h, RK45.A, RK45.B, RK45.C, K)
    err_scipy = K.T @ RK45.E * h
    yt = torch.tensor(y, dtype=F64)
    y1, f1, err, k = tde._runge_kutta_step(_nonlin, yt, _nonlin(0.0, yt), 0.0, h)
    assert np.abs(y1.numpy() - y_new).max() < 1e-15
    assert np.abs(f1.numpy() - f_new).max() < 1e-15
    assert np.abs(err.numpy() + (2.0 / 3.0) * err_scipy).max() < 1e-17 + 1e-12 * np.abs(err_scipy).max()
    for j in range(7):
        assert np.abs(k[j].numpy() - K[j]).max() < 1e-15
    # dense output: torchdiffeq's quartic through (y0, f0, y_mid, y1, f1) IS scipy's RkDenseOutput polynomial
    coeffs = tde._dopri5_interp_fit(yt, y1, k, h)
    Q = K.T @ RK45.P
    dense = scipy_rk.RkDenseOutput(0.0, h, y, Q)
    for x in (0.0, 0.1, 0.5, 0.77, 1.0):
        got = tde._interp_evaluate(coeffs, torch.tensor(0.0, dtype=F64), torch.tensor(h, dtype=F64),
                                   torch.tensor(x * h, dtype=F64))
        assert np.abs(got.numpy() - dense(x * h)).max() < 2e-14, x


@pytest.mark.parametrize("rtol,atol", [(1e-3, 1e-4), (1e-4, 1e-5), (1e-7, 1e-9)])
def test_initial_step_equals_scipy_select_initial_step(rtol, atol):
    """misc._select_initial_step (called with order = 5 - 1 = 4) against scipy.integrate._ivp.common.select_initial_step
    (called by RK45 with its error_estimator_order = 4): same Hairer heuristic, same RMS norm -> same number."""
    y0 = np.array([0.7, -0.4, 1.5, 0.01])
    fun_np = lambda t, y: np.array([math.sin(y[1]) + 0.3 * y[0], math.cos(y[0]) * y[2], -y[2] + y[3], y[0] * y[3]])   # noqa
    fun_t = lambda t, y: torch.stack([torch.sin(y[1]) + 0.3 * y[0], torch.cos(y[0]) * y[2], -y[2] + y[3], y[0] * y[3]])  # noqa
    try:
        want = scipy_common.select_initial_step(fun_np, 0.0, y0, 1e9, np.inf, fun_np(0.0, y0), 1, 4, rtol, atol)
    except TypeError:                       # scipy < 1.11 has no t_bound / max_step arguments
        want = scipy_common.select_initial_step(fun_np, 0.0, y0, fun_np(0.0, y0), 1, 4, rtol, atol)
    yt = torch.tensor(y0, dtype=F64)
    log = tde.StepLog()
    got = tde._select_initial_step(fun_t, torch.tensor(0.0, dtype=F64), yt, 4, rtol, atol, fun_t(0.0, yt), log)
    assert log.n_feval == 1
    assert abs(float(got) - want) <= 1e-13 * want, (float(got), want)


@pytest.mark.parametrize("rtol,atol", [(1e-3, 1e-4), (1e-5, 1e-7), (1e-8, 1e-10)])
def test_dopri5_solve_agrees_with_scipy_rk45_to_tolerance(rtol, atol):
    y0 = [0.7, -0.4]
    t = np.linspace(0.0, 4.0, 21)
    ref = solve_ivp(_nonlin_np, (0.0, 4.0), y0, method="DOP853", rtol=1e-13, atol=1e-14, t_eval=t).y.T
    sp = solve_ivp(_nonlin_np, (0.0, 4.0), y0, method="RK45", rtol=rtol, atol=atol, t_eval=t)
    log = tde.StepLog()
    sol = tde.odeint(_nonlin, torch.tensor(y0, dtype=F64), torch.tensor(t, dtype=F64), rtol=rtol, atol=atol,
                     method="dopri5", log=log).numpy()
    e_ours, e_scipy = np.abs(sol - ref).max(), np.abs(sp.y.T - ref).max()
    assert e_ours < 50 * rtol and e_ours < 20 * max(e_scipy, atol), (e_ours, e_scipy)
    # the first attempt is the same in both (same initial step); a comparable number of f-evals after that
    assert log.n_feval == 2 + 6 * len(log.attempts)
    assert 0.3 * sp.nfev <= log.n_feval <= 3 * sp.nfev, (log.n_feval, sp.nfev)
    # steps are not clipped to the output times: fewer accepted steps than output intervals at the loose tolerance
    if rtol >= 1e-3:
        assert log.n_accept < len(t) - 1


def test_dopri5_first_attempt_uses_the_initial_step_and_outputs_interpolate():
    log = tde.StepLog()
    t = torch.tensor([0.0, 0.01, 0.02, 3.0], dtype=F64)
    y0 = torch.tensor([0.7, -0.4], dtype=F64)
    sol = tde.odeint(_nonlin, y0, t, rtol=1e-6, atol=1e-8, method="dopri5", log=log)
    assert log.attempts[0][0] == 0.0 and log.attempts[0][1] == log.dt0
    # t = 0.01 and 0.02 both lie inside the first accepted step (dense output, no clipping)
    first_acc = next(a for a in log.attempts if a[3])
    assert first_acc[0] + first_acc[1] > 0.02
    ref = solve_ivp(_nonlin_np, (0.0, 3.0), [0.7, -0.4], method="DOP853", rtol=1e-13, atol=1e-14, t_eval=t.numpy()).y.T
    assert np.abs(sol.numpy() - ref).max() < 1e-5


# ------------------------------------------------------------------------------------------------ controller rules
def test_optimal_step_size_rules():
    """misc._optimal_step_size: err == 0 -> x10; err < 1 -> never shrink (dfactor = 1); err >= 1 -> shrink by
    0.9 * err^(-1/5), at most 5x; growth capped at 10x."""
    dt = torch.tensor(0.1, dtype=F64)
    assert float(tde._optimal_step_size(dt, 0.0)) == pytest.approx(1.0)
    assert float(tde._optimal_step_size(dt, 1e-12)) == pytest.approx(1.0)                 # ifactor cap
    assert float(tde._optimal_step_size(dt, 0.9)) == pytest.approx(0.1)                   # 0.9*0.9^-0.2 < 1 -> dfactor 1
    assert float(tde._optimal_step_size(dt, 0.5)) == pytest.approx(0.1 * 0.9 * 0.5 ** -0.2)
    assert float(tde._optimal_step_size(dt, 2.0)) == pytest.approx(0.1 * 0.9 * 2.0 ** -0.2)
    assert float(tde._optimal_step_size(dt, 1e9)) == pytest.approx(0.02)                  # dfactor floor
    assert float(tde._optimal_step_size(dt, 1.0)) == pytest.approx(0.09)                  # err == 1 is accepted and shrinks


def test_rejected_attempts_keep_state_and_shrink():
    """A stiff-ish linear problem with a far too large first step: rejected attempts restart from the same t0 with a smaller
    dt, accepted ones advance t0 by exactly dt, and t0 never moves backwards."""
    log = tde.StepLog()
    lam = -400.0
    t = torch.tensor([0.0, 1.0], dtype=F64)
    sol = tde.odeint(lambda t_, y: lam * y + 3.0, torch.tensor([1.0, 2.0], dtype=F64), t, rtol=1e-3, atol=1e-6,
                     method="dopri5", log=log)
    assert log.n_reject > 0
    for a, b in zip(log.attempts[:-1], log.attempts[1:]):
        if a[3]:
            assert b[0] == pytest.approx(a[0] + a[1], abs=1e-15) and b[1] >= a[1] * (1 - 1e-12)
        else:
            assert b[0] == a[0] and 0.2 * a[1] * (1 - 1e-12) <= b[1] < a[1]
    exact = 3.0 / 400.0 + (torch.tensor([1.0, 2.0], dtype=F64) - 3.0 / 400.0) * math.exp(lam)
    assert float((sol[-1] - exact).abs().max()) < 1e-4


def test_rms_norm_is_over_the_whole_batch_tensor():
    """one step size for the whole batch -- adding a sample with large error changes everyone's steps."""
    f = lambda t_, y: -y * torch.tensor([[1.0], [60.0]], dtype=F64)      # noqa: E731
    t = torch.tensor([0.0, 1.0], dtype=F64)
    la, lb = tde.StepLog(), tde.StepLog()
    tde.odeint(lambda t_, y: -y, torch.ones(1, 3, dtype=F64), t, rtol=1e-5, atol=1e-7, method="dopri5", log=la)
    tde.odeint(f, torch.ones(2, 3, dtype=F64), t, rtol=1e-5, atol=1e-7, method="dopri5", log=lb)
    assert len(lb.attempts) > len(la.attempts)


# ------------------------------------------------------------------------------------------------ degenerate / API cases
def test_single_time_point_returns_y0_and_memory_variant_doubles():
    """odeint with len(t) == 1 performs no integration; modules/DiffEqSolver.py:30-42 (memory=True) therefore
    yields y_i = 2^i * z0."""
    y0 = torch.randn(2, 3, 4, 4)
    calls = []

    def f(t_, y):
        calls.append(1)
        return y

    for method in ("euler", "rk4", "dopri5"):
        out = tde.odeint(f, y0, torch.tensor([0.3], dtype=F64), method=method)
        assert out.shape == (1,) + y0.shape and torch.equal(out[0], y0)
    assert not calls
    h = y0
    for i in range(1, 4):
        h = h + tde.odeint(f, h, torch.tensor([0.1 * i], dtype=F64), method="dopri5")[0]
        assert torch.equal(h, y0 * 2 ** i)


def test_state_dtype_fp32_time_fp64():
    """fp32 state, fp64 time: the solution stays fp32, dt in the log keeps fp64 resolution."""
    log = tde.StepLog()
    t = torch.tensor(np.arange(10, 20) / 20)                    # helpers/utils.py:118-120
    sol = tde.odeint(lambda t_, y: -y, torch.ones(4, dtype=torch.float32), t, rtol=1e-3, atol=1e-4, method="dopri5", log=log)
    assert sol.dtype == torch.float32 and t.dtype == F64
    assert float((sol[-1] - math.exp(-0.45)).abs().max()) < 1e-3
    assert any(a[1] != float(np.float32(a[1])) for a in log.attempts)


def test_unsupported_inputs_raise():
    y0 = torch.ones(2)
    with pytest.raises(AssertionError):
        tde.odeint(lambda t_, y: y, y0, torch.tensor([[0.0, 1.0]], dtype=F64), method="euler")
    with pytest.raises(ValueError):
        tde.odeint(lambda t_, y: y, y0, torch.tensor([0.0, 1.0], dtype=F64), method="adams")
