"""Host-side logic of the product (solver loops, speculative dopri5 chunks, hand-written backward replay, layout
handling, drop-in surface) checked on CPU against the oracle, with the CUDA backend replaced by the torch-CPU
emulation of tests/fake_backend.py.  The kernels themselves are checked by the `-m gpu` tests."""
import pytest
import torch

import ode_rl_b200 as odex
from oracle import refmodel, torchdiffeq_021 as tde
from tests.util import make_ode_func, rel_err, timesteps


def _sd(m):
    return {k: v.detach().clone() for k, v in m.state_dict().items()}


@pytest.mark.parametrize("method", ["euler", "rk4"])
def test_fixed_grid_forward_and_grad(fake_backend, method):
    f = make_ode_func(odex, C=32, n_units=32)
    solver = odex.DiffEqSolver(f, method)
    _, tp = timesteps(3, 4)
    torch.manual_seed(1)
    z0 = torch.randn(2, 32, 8, 16, requires_grad=True)
    sol = solver(z0, tp)
    assert sol.shape == (4, 2, 32, 8, 16)
    w = torch.randn_like(sol)
    (sol * w).sum().backward()
    got = {n: p.grad.clone() for n, p in solver.named_parameters()}
    gz = z0.grad.clone()
    # oracle: autograd through the restated odeint over the same weights
    sd = _sd(solver)
    ps = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z1 = z0.detach().clone().requires_grad_(True)
    ref = refmodel.diffeq_solver(ps, z1, tp, method)
    (ref * w).sum().backward()
    assert rel_err(sol, ref) < 1e-5
    assert rel_err(gz, z1.grad) < 1e-4
    for n, g in got.items():
        assert rel_err(g, ps[n].grad) < 1e-4, n


def test_dopri5_forward_matches_oracle_step_for_step(fake_backend):
    f = make_ode_func(odex, C=32, n_units=32, scale=3.0)
    solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
    solver._chunk_hint = 2                         # force several speculative chunks
    _, tp = timesteps(10, 10)
    torch.manual_seed(2)
    z0 = torch.randn(2, 32, 8, 16)
    with torch.no_grad():
        sol = solver(z0, tp)
    log = tde.StepLog()
    ref = refmodel.diffeq_solver(_sd(solver), z0, tp, "dopri5", 1e-3, 1e-4, log=log)
    st = solver.last_stats
    assert (st.n_accept, st.n_reject) == (log.n_accept, log.n_reject)
    assert st.n_feval == log.n_feval
    assert st.host_syncs >= 2
    for a, b in zip(st.attempts, log.attempts):
        assert abs(a[1] - b[1]) <= 2e-5 * abs(b[1]) and bool(a[3]) == b[3]
    assert rel_err(sol, ref) < 1e-5
    # second call: the chunk hint makes it a single host sync
    with torch.no_grad():
        solver(z0, tp)
    assert solver.last_stats.host_syncs == 1


@pytest.mark.parametrize("nonlinear", ["tanh", "relu"])
def test_dopri5_gradients(fake_backend, nonlinear):
    """Gradient through accepted steps + dense output.  With a smooth activation the hand-written replay matches
    autograd through the oracle to fp32 rounding; with ReLU a pre-activation within rounding noise of 0 can take
    the other branch (both are valid sub-gradients), so that case is held to an L2 bound instead of max-abs."""
    import torch.nn.functional as F

    f = make_ode_func(odex, C=32, n_units=32, scale=2.0, nonlinear=nonlinear)
    solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
    _, tp = timesteps(4, 6)
    torch.manual_seed(3)
    z0 = torch.randn(2, 32, 8, 16, requires_grad=True)
    sol = solver(z0, tp)
    w = torch.randn_like(sol)
    (sol * w).sum().backward()
    ps = {k: v.clone().requires_grad_(True) for k, v in _sd(solver).items()}
    z1 = z0.detach().clone().requires_grad_(True)

    def fn(t, y):
        for n, i in enumerate((0, 2, 4, 6, 8)):
            if n:
                y = F.relu(y) if nonlinear == "relu" else torch.tanh(y)
            y = F.conv2d(y, ps["ode_func.gradient_net.%d.weight" % i], ps["ode_func.gradient_net.%d.bias" % i], padding=1)
        return y

    ref = tde.odeint(fn, z1, tp, rtol=1e-3, atol=1e-4, method="dopri5")
    (ref * w).sum().backward()
    assert rel_err(sol, ref) < 1e-5
    from tests.util import rel_l2
    if nonlinear == "tanh":
        assert rel_err(z0.grad, z1.grad) < 1e-5
        for n, p in solver.named_parameters():
            assert rel_err(p.grad, ps[n].grad) < 1e-5, n
    else:
        assert rel_l2(z0.grad, z1.grad) < 1e-3
        for n, p in solver.named_parameters():
            assert rel_l2(p.grad, ps[n].grad) < 1e-3, n


@pytest.mark.parametrize("C,hw", [(32, (8, 16)), (64, (16, 16))], ids=["c32", "fused_chain_shape"])
def test_encoder_recurrence_forward_and_grad(fake_backend, C, hw):
    """(64, 16x16): f runs as the fused chain with per-image maxima in a [L, 2, T, B] arena, and the weight gradients merged over
    the T iterations take them (the fake backend checks every pair against the operands it is handed)."""
    torch.manual_seed(0)
    f = make_ode_func(odex, C=C, n_units=C)
    cell = odex.ODEConvGRUCell(f, None, hw, C, device=torch.device("cpu"))
    obs, _ = timesteps(5, 5)
    torch.manual_seed(4)
    x = torch.randn(5, 2, C, hw[0], hw[1], requires_grad=True)
    n_scaled = fake_backend.wgrad_scaled
    mean, std = cell(x, obs)
    w1, w2 = torch.randn_like(mean), torch.randn_like(std)
    ((mean * w1).sum() + (std * w2).sum()).backward()
    ps = {k: v.clone().requires_grad_(True) for k, v in _sd(cell).items()}
    x1 = x.detach().clone().requires_grad_(True)
    rm, rs = refmodel.encoder_z0(ps, x1, obs)
    ((rm * w1).sum() + (rs * w2).sum()).backward()
    assert rel_err(mean, rm) < 1e-5 and rel_err(std, rs) < 1e-5
    assert rel_err(x.grad, x1.grad) < 1e-4
    for n, p in cell.named_parameters():
        assert rel_err(p.grad, ps[n].grad) < 2e-4, n
    assert (fake_backend.wgrad_scaled > n_scaled) == (C == 64)
    yi, lat = cell.run_ode_conv_gru(x.detach(), obs)
    ry, rl = refmodel.run_ode_conv_gru(_sd(cell), x.detach(), obs)
    assert lat.shape == rl.shape and rel_err(lat, rl) < 1e-5 and rel_err(yi, ry) < 1e-5


def test_convgru_cell_multi_step(fake_backend):
    torch.manual_seed(0)
    cell = odex.ConvGRUCell((8, 16), 32, 32, 5)
    x = torch.randn(3, 2, 32, 8, 16)
    h = torch.randn(2, 32, 8, 16)
    out, hn = cell(x, h, seq_len=3)
    sd = _sd(cell)
    hh = h
    for i in range(3):
        hh = refmodel.convgru_step(sd, x[i], hh)
    assert out.shape == (3, 2, 32, 8, 16)
    assert rel_err(hn, hh) < 1e-5 and rel_err(out[-1], hh) < 1e-5


def test_memory_mode_and_errors(fake_backend):
    f = make_ode_func(odex, C=32, n_units=32)
    _, tp = timesteps(2, 3)
    z0 = torch.randn(1, 32, 8, 16)
    y = odex.DiffEqSolver(f, "euler", memory=True)(z0, tp)     # degenerate 2^i * z0
    assert y.shape == (1, 3, 32, 8, 16)
    assert rel_err(y[:, 2], 8 * z0) < 1e-6
    with pytest.raises(ValueError):
        odex.DiffEqSolver(f, "adams")(z0, tp)
    cell = odex.ODEConvGRUCell(f, None, (8, 16), 32)
    with pytest.raises(AssertionError):
        cell(torch.randn(4, 1, 32, 8, 16), tp)


def test_product_has_no_cpu_path():
    """Without the fake backend the product refuses CPU tensors (no silent fallback)."""
    f = make_ode_func(odex, C=32, n_units=32)
    _, tp = timesteps(2, 3)
    with pytest.raises(odex.OdexError):
        odex.DiffEqSolver(f, "euler")(torch.randn(1, 32, 8, 16), tp)


def test_c_abi_library_exports_every_header_symbol():
    """libodex.so loads on a CPU-only box (static cudart) and exports exactly what include/odex.h declares; the
    ctypes prototype table covers every one of them.  No compute call is made."""
    import ctypes
    import os
    import re

    from ode_rl_b200 import _lib

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "odex.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(odex_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    l = _lib.lib()
    for name in declared:
        assert hasattr(l, name), name
    assert l.odex_version() == _lib.ODEX_VERSION
    assert ctypes.sizeof(_lib.DopriState) <= _lib.STATE_HEAD_BYTES
    assert l.odex_dopri_state_bytes() > _lib.STATE_HEAD_BYTES


def test_grad_sink_merges_consecutive_slices(fake_backend):
    """GradSink.add_wgrad while deferred: requests whose (x, dy) are consecutive dense slices of one buffer become ONE
    launch over n*B images; anything else stays a launch of its own; the accumulated gradient equals the eager one."""
    import torch.nn as nn
    from ode_rl_b200.ops import GradSink, _slice_runs

    torch.manual_seed(3)
    conv = nn.Conv2d(8, 8, 3, padding=1)
    xa, da = torch.randn(4, 2, 6, 6, 8), torch.randn(4, 2, 6, 6, 8)          # [s, B, H, W, C] arenas
    lone_x, lone_d = torch.randn(2, 6, 6, 8), torch.randn(2, 6, 6, 8)
    items = [(xa[2], da[2]), (lone_x, lone_d), (xa[0], da[0]), (xa[1], da[1]), (xa[3], da[1])]   # last: dy not consecutive
    runs = _slice_runs(items)
    sizes = sorted(r[0].shape[0] for r in runs)
    assert sizes == [2, 2, 6]                                                # 3 slices merged, two singles
    merged = [r for r in runs if r[0].shape[0] == 6][0]
    assert torch.equal(merged[0], xa[:3].reshape(6, 6, 6, 8)) and torch.equal(merged[1], da[:3].reshape(6, 6, 6, 8))

    def total(deferred):
        sink = GradSink()
        sink.deferred = deferred
        n0 = fake_backend.launches
        for x, d in items:
            sink.add_wgrad(conv, x, d)
        gw, gb = sink.result([conv.weight, conv.bias])
        return gw, gb, fake_backend.launches - n0

    gw0, gb0, l0 = total(False)
    gw1, gb1, l1 = total(True)
    assert l1 == l0 - 2                                                      # 5 launches -> 3 (+ the same unpack)
    assert rel_err(gw1, gw0) < 1e-6 and rel_err(gb1, gb0) < 1e-6


def test_empty_batch_returns_empty_results_with_zero_grads(fake_backend):
    """A ragged shard may hold no sample (parallel.shard_range(2, r, 4)): the modules return empty tensors of the right
    shape without launching anything, and backward() gives the parameters zero gradients (what torch's conv does for a
    0-sample batch).  A sharded dopri5 solver refuses instead: every rank must contribute to the shared step control."""
    f = make_ode_func(odex, C=32, n_units=32)
    cell = odex.ODEConvGRUCell(f, None, (8, 16), 32)
    obs, tp = timesteps(3, 3)
    n0 = fake_backend.launches
    for method in ("euler", "rk4", "dopri5"):
        solver = odex.DiffEqSolver(f, method)
        z0 = torch.zeros(0, 32, 8, 16, requires_grad=True)
        sol = solver(z0, tp)
        assert sol.shape == (3, 0, 32, 8, 16)
        sol.sum().backward()
        assert all(p.grad is not None and float(p.grad.abs().sum()) == 0.0 for p in solver.parameters())
    mean, std = cell(torch.zeros(3, 0, 32, 8, 16), obs)
    assert mean.shape == (0, 32, 8, 16) and std.shape == (0, 32, 8, 16)
    assert f(None, torch.zeros(0, 32, 8, 16)).shape == (0, 32, 8, 16)
    assert fake_backend.launches == n0
    sharded = odex.DiffEqSolver(f, "dopri5")
    sharded.reduce_fn = lambda t: t
    with pytest.raises(ValueError):
        sharded(torch.zeros(0, 32, 8, 16), tp)


def test_conv_transpose4x4s2_phase_mapping(fake_backend):
    """ConvTranspose2d(k=4, s=2, p=1) written as one 3x3 convolution to 4*Cout phase channels + depth-to-space
    (modules.conv_transpose4x4s2): forward and all three gradients against torch's own operator."""
    import torch.nn as nn
    import torch.nn.functional as F
    from ode_rl_b200.modules import _ConvT4x4s2Fn, phases_to_nchw

    torch.manual_seed(2)
    m = nn.ConvTranspose2d(64, 32, 4, 2, 1)
    x = torch.randn(3, 64, 8, 16, requires_grad=True)
    y = phases_to_nchw(_ConvT4x4s2Fn.apply(x, m.weight, m.bias))
    ref = F.conv_transpose2d(x, m.weight, m.bias, stride=2, padding=1)
    assert y.shape == ref.shape == (3, 32, 16, 32)
    assert rel_err(y, ref) < 1e-5
    w = torch.randn_like(ref)
    gx, gw, gb = torch.autograd.grad((y * w).sum(), [x, m.weight, m.bias])
    rx, rw, rb = torch.autograd.grad((ref * w).sum(), [x, m.weight, m.bias])
    assert rel_err(gx, rx) < 1e-5 and rel_err(gw, rw) < 1e-5 and rel_err(gb, rb) < 1e-5


def test_conv3x3s2_space_to_depth_mapping(fake_backend):
    """Conv2d(k=3, s=2, p=1) written as a 3x3 stride-1 convolution over space-to-depth cells (modules.conv3x3s2):
    forward and all three gradients against torch's own operator."""
    import torch.nn as nn
    import torch.nn.functional as F
    from ode_rl_b200.modules import _Conv3x3s2Fn

    torch.manual_seed(2)
    m = nn.Conv2d(16, 64, 3, 2, 1)
    x = torch.randn(3, 16, 16, 32, requires_grad=True)
    y = _Conv3x3s2Fn.apply(x, m.weight, m.bias)
    ref = F.conv2d(x, m.weight, m.bias, stride=2, padding=1)
    assert y.shape == ref.shape == (3, 64, 8, 16)
    assert rel_err(y, ref) < 1e-5
    w = torch.randn_like(ref)
    gx, gw, gb = torch.autograd.grad((y * w).sum(), [x, m.weight, m.bias])
    rx, rw, rb = torch.autograd.grad((ref * w).sum(), [x, m.weight, m.bias])
    assert rel_err(gx, rx) < 1e-5 and rel_err(gw, rw) < 1e-5 and rel_err(gb, rb) < 1e-5


def test_decode_tail_function(fake_backend):
    """modules.decode_tail (LeakyReLU + ConvTranspose2d(32, 1, 4, 2, 1) on the phase-blocked tensor) wired through autograd:
    output and the three gradients against the torch ops on the depth-to-space tensor."""
    import torch.nn as nn
    from ode_rl_b200.modules import decode_tail, decode_tail_supported, phases_to_nchw

    torch.manual_seed(4)
    act, m = nn.LeakyReLU(0.2), nn.ConvTranspose2d(32, 1, 4, 2, 1)
    P = torch.randn(2, 16, 16, 128, requires_grad=True)
    assert decode_tail_supported(P, act, m)
    y = decode_tail(P, act, m)
    ref = m(act(phases_to_nchw(P)))
    assert y.shape == ref.shape == (2, 1, 64, 64) and rel_err(y, ref) < 1e-6
    w = torch.randn_like(ref)
    g = torch.autograd.grad((y * w).sum(), [P, m.weight, m.bias])
    r = torch.autograd.grad((ref * w).sum(), [P, m.weight, m.bias])
    assert all(rel_err(a, b) < 1e-5 for a, b in zip(g, r))


def test_upsample2x_autograd_wiring(fake_backend):
    """modules.upsample2x (forward kernel + adjoint kernel behind one autograd node) against torch's interpolate."""
    import torch.nn.functional as F
    from ode_rl_b200 import modules

    torch.manual_seed(3)
    x = torch.randn(2, 8, 5, 6, requires_grad=True)
    fake_backend.require = lambda t: None
    y = modules._Upsample2xFn.apply(x)
    ref = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=False)
    assert rel_err(y, ref) < 1e-6
    w = torch.randn_like(ref)
    (g,) = torch.autograd.grad((y * w).sum(), [x])
    (r,) = torch.autograd.grad((ref * w).sum(), [x])
    assert rel_err(g, r) < 1e-6


def _dopri_error_cases(odex, device):
    """torchdiffeq's asserts (rk_common._adaptive_step; SURVEY.md 8b 'Errors') raised once per forward from the device-side
    sticky status word: non-finite state (surfacing as torchdiffeq does, through the dt check), max_num_steps."""
    f = make_ode_func(odex, C=32, n_units=32, scale=4.0).to(device)
    _, tp = timesteps(4, 4)
    tp = tp.to(device)
    torch.manual_seed(1)
    z0 = torch.randn(2, 32, 8, 16, device=device)
    solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
    bad = z0.clone()
    bad[1, 3, 2, 5] = float("nan")
    with torch.no_grad():
        # a NaN in y0 makes the initial step NaN, and torchdiffeq's FIRST assert in _adaptive_step is `t0 + dt > t0`:
        # the oracle raises 'underflow in dt nan' for this input and so must the product
        with pytest.raises(AssertionError, match="underflow in dt nan"):
            solver(bad, tp)
        tight = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-9, odeint_atol=1e-10)
        tight.max_attempts = 3
        with pytest.raises(AssertionError, match="max_num_steps"):
            tight(z0, tp)
        sol = solver(z0, tp)                    # the sticky flags do not leak into the next forward
    assert bool(torch.isfinite(sol).all()) and solver.last_stats.status == 0


def test_dopri5_error_behaviour(fake_backend):
    _dopri_error_cases(odex, torch.device("cpu"))


def test_encode_stem_autograd_wiring(fake_backend):
    """modules.encode_stem_then_conv (stem kernel -> space-to-depth cells -> strided convolution on the cells, one autograd
    node each) against torch's layers: output and the four parameter gradients."""
    import torch.nn as nn
    import torch.nn.functional as F
    from ode_rl_b200 import modules

    fake_backend.require = lambda t: None
    torch.manual_seed(0)
    c0, c2 = nn.Conv2d(1, 16, 3, 2, 1), nn.Conv2d(16, 64, 3, 2, 1)
    x = torch.rand(3, 1, 64, 64)
    y = modules.encode_stem_then_conv(x, c0, nn.LeakyReLU(0.2), c2)
    ref = c2(F.leaky_relu(c0(x), 0.2))
    assert rel_err(y, ref) < 1e-6
    w = torch.randn_like(ref)
    ps = [c0.weight, c0.bias, c2.weight, c2.bias]
    got = torch.autograd.grad((torch.tanh(y) * w).sum(), ps)
    exp = torch.autograd.grad((torch.tanh(ref) * w).sum(), ps)
    for g, r in zip(got, exp):
        assert rel_err(g, r) < 1e-5


# ---- round-2 robustness items (VERDICT r1 weak #12 / missing #7) ------------------------------------------
def _robustness_cases(odex, device, C=32, hw=(8, 16)):
    """Runs on the fake backend here and on the CUDA kernels in tests/test_gpu_parity.py (C = 64, 16 x 16: kernel shapes)."""
    import gc
    import numpy as np

    f = make_ode_func(odex, C=C, n_units=C, scale=2.0).to(device)
    cell = odex.ODEConvGRUCell(f, None, hw, C).to(device)
    torch.manual_seed(5)
    x = torch.randn(4, 2, C, *hw, device=device)
    sd = {k: v.detach().cpu().clone() for k, v in cell.state_dict().items()}

    # (1) the encoder step sizes follow the VALUES of `timesteps`: a second tensor of the same shape (allocated where the
    #     first one lived after it is freed) must not hit a stale cache (ADVICE: modules.py:886)
    for denom in (8.0, 12.0, 8.0):
        ts = torch.tensor(np.arange(4) / denom).to(device)
        with torch.no_grad():
            m, _ = cell(x, ts)
        rm, _ = refmodel.encoder_z0(sd, x.cpu(), ts.cpu())
        assert rel_err(m, rm) < 1e-5, denom
        del ts
        gc.collect()
    irregular = torch.tensor([0.0, 0.03, 0.2, 0.21], dtype=torch.float64)
    with torch.no_grad():
        m, _ = cell(x, irregular.to(device))
    assert rel_err(m, refmodel.encoder_z0(sd, x.cpu(), irregular)[0]) < 1e-5

    # (2) NaN in the recurrence -> AssertionError from the sticky device word, once per forward; the word is cleared
    bad = x.clone()
    bad[2, 1, 3, 4, 5] = float("nan")
    with torch.no_grad():
        with pytest.raises(AssertionError, match="NaN"):
            cell(bad, irregular.to(device))
        cell(x, irregular.to(device))

    # (3) a second backward through the same graph fails loudly instead of with an AttributeError deep inside
    solver = odex.DiffEqSolver(f, "rk4")
    z0 = torch.randn(2, C, *hw, device=device, requires_grad=True)
    _, tp = timesteps(3, 3)
    for make in (lambda: solver(z0, tp.to(device)), lambda: cell(x.clone().requires_grad_(True), irregular.to(device))[0],
                 lambda: odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)(z0, tp.to(device))):
        out = make()
        loss = out.square().sum()
        loss.backward(retain_graph=True)
        with pytest.raises(RuntimeError, match="already back-propagated"):
            loss.backward()

    # (4) time grids: unsorted t is an error (torchdiffeq asserts), for dopri5 from the host copy the forward reads anyway;
    #     a decreasing grid is exact for the fixed-grid solvers (negative dt)
    unsorted = torch.tensor([0.5, 0.7, 0.6, 0.8], dtype=torch.float64)
    with torch.no_grad():
        with pytest.raises(AssertionError, match="strictly increasing"):
            odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)(z0.detach(), unsorted.to(device))
        with pytest.raises(AssertionError, match="strictly increasing or decreasing"):
            odex.DiffEqSolver(f, "euler")(z0.detach(), unsorted)
        dec = torch.tensor([0.8, 0.7, 0.55, 0.5], dtype=torch.float64)
        sol = odex.DiffEqSolver(f, "rk4")(z0.detach(), dec.to(device))
    fsd = {k: v.detach().cpu().clone() for k, v in solver.state_dict().items()}
    y, ref = z0.detach().cpu(), [z0.detach().cpu()]
    for a, b in zip(dec[:-1], dec[1:]):
        y = y + tde._rk4_alt_step(lambda t_, v: refmodel.ode_func({k[len("ode_func."):]: w for k, w in fsd.items()}, v), a, b - a, y)
        ref.append(y)
    assert rel_err(sol, torch.stack(ref)) < 1e-5

    # (5) more accepted steps than backward checkpoints: the device stops there and the host says so (no silent drop)
    tight = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-7, odeint_atol=1e-9)
    tight.save_budget_bytes = 1                          # -> the minimum capacity of 4 steps
    with pytest.raises(RuntimeError, match="checkpoint capacity"):
        tight(z0, timesteps(10, 10)[1].to(device))
    with torch.no_grad():                                # without a backward there is no capacity to exceed
        tight(z0.detach(), timesteps(10, 10)[1].to(device))
    assert tight.last_stats.n_accept > 4 and tight.last_stats.host_syncs >= 1


def test_robustness_cases(fake_backend):
    _robustness_cases(odex, torch.device("cpu"))


def test_dopri5_one_read_back_carries_state_times_and_log(fake_backend):
    """The forward's single device->host copy holds the controller header, the output times (the backward needs them) and
    the attempt log: no other host read on the training path."""
    f = make_ode_func(odex, C=32, n_units=32, scale=3.0)
    solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
    _, tp = timesteps(10, 10)
    z0 = torch.randn(2, 32, 8, 16, requires_grad=True)
    solver(z0, tp)                                        # learns the chunk size
    calls = []
    orig = torch.Tensor.cpu

    def counting_cpu(self, *a, **k):
        calls.append(tuple(self.shape))
        return orig(self, *a, **k)

    torch.Tensor.cpu = counting_cpu
    try:
        sol = solver(z0, tp)
    finally:
        torch.Tensor.cpu = orig
    st = solver.last_stats
    assert len(calls) == 1 and st.host_syncs == 1
    assert st.t_host == tp.tolist() and len(st.attempts) == st.n_attempt
    sol.sum().backward()


@pytest.mark.parametrize("rtol,atol", [(1e-3, 1e-4), (1e-5, 1e-6)])
def test_dt0_detach_deviation_is_of_the_order_of_rtol(rtol, atol):
    """the product (and the oracle's default) treat the initial step dt0 as a constant in the backward;
    torchdiffeq 0.2.1 does not detach _select_initial_step.  dt0 only enters through the FIRST attempt's
    step size, and the converged solution depends on step sizes at the level of the local error, so the extra gradient path
    is O(rtol) relative to the gradient -- the same size as the solver's own discretisation error.  Measured on the oracle,
    both ways (1.7e-3 at rtol 1e-3)."""
    torch.manual_seed(0)
    f = make_ode_func(odex, C=16, n_units=16, scale=3.0, nonlinear="tanh")
    sd = _sd(odex.DiffEqSolver(f, "dopri5"))
    _, tp = timesteps(5, 5)
    z0 = torch.randn(2, 16, 8, 8)
    grads = []
    for detach in (True, False):
        ps = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        z = z0.clone().requires_grad_(True)
        fn = lambda t_, y: refmodel.ode_func({k[len("ode_func."):]: w for k, w in ps.items()}, y, "tanh")   # noqa: E731
        sol = tde.odeint(fn, z, tp, rtol=rtol, atol=atol, method="dopri5", detach_dt0=detach)
        sol[-1].square().sum().backward()
        grads.append(torch.cat([z.grad.flatten()] + [p.grad.flatten() for p in ps.values() if p.grad is not None]))
    gap = float((grads[0] - grads[1]).norm() / grads[1].norm())
    print("dt0 gradient path at rtol %g: relative gap %.2e" % (rtol, gap))
    assert gap < 5 * rtol


def test_bn_relu_and_warp_compose_autograd_wiring(fake_backend):
    """modules.bn_relu (BatchNorm2d batch statistics + ReLU, running-statistics update) and modules.warp_compose (one step of the
    VidODE flow warp + mask composition) against torch's own modules: values, gradients, module buffers."""
    import torch.nn as nn
    import torch.nn.functional as F
    from ode_rl_b200 import modules

    fake_backend.require = lambda t: None
    torch.manual_seed(0)
    bn_a, bn_b = nn.BatchNorm2d(8), nn.BatchNorm2d(8)
    with torch.no_grad():
        bn_a.weight.uniform_(0.5, 1.5); bn_a.bias.normal_(0, 0.2)
    bn_b.load_state_dict(bn_a.state_dict())
    x = torch.randn(3, 8, 6, 5).contiguous(memory_format=torch.channels_last)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya = modules.bn_relu(xa, bn_a)
    yb = F.relu(bn_b(xb))
    assert rel_err(ya, yb) < 1e-6
    w = torch.randn_like(yb)
    (ya * w).sum().backward()
    (yb * w).sum().backward()
    assert rel_err(xa.grad, xb.grad) < 1e-5
    assert rel_err(bn_a.weight.grad, bn_b.weight.grad) < 1e-5 and rel_err(bn_a.bias.grad, bn_b.bias.grad) < 1e-5
    assert rel_err(bn_a.running_mean, bn_b.running_mean) < 1e-6 and rel_err(bn_a.running_var, bn_b.running_var) < 1e-6
    assert int(bn_a.num_batches_tracked) == 1
    # warp + compose
    b, c, h, wd = 2, 3, 8, 10
    prev = torch.rand(b, c, h, wd)
    head = torch.randn(b, 8, h, wd)[:, :c + 3]                 # a channel slice of a wider tensor, like the padded decoder head
    head = head * torch.tensor([3.0, 3.0, 1, 1, 1, 1]).view(1, 6, 1, 1)
    lin_x, lin_y = torch.linspace(-1.0, 1.0, wd), torch.linspace(-1.0, 1.0, h)
    pa, ha = prev.clone().requires_grad_(True), head.clone().requires_grad_(True)
    warped, pred = modules.warp_compose(pa, ha, lin_x, lin_y)
    pb, hb = prev.clone().requires_grad_(True), head.clone().requires_grad_(True)
    gx = lin_x.view(1, 1, wd, 1).expand(b, h, -1, -1)
    gy = lin_y.view(1, h, 1, 1).expand(b, -1, wd, -1)
    grid = torch.cat([gx, gy], 3)
    fl = torch.cat([hb[:, 0:1] / ((wd - 1.0) / 2.0), hb[:, 1:2] / ((h - 1.0) / 2.0)], dim=1)
    rw = F.grid_sample(pb, grid + fl.permute(0, 2, 3, 1), padding_mode="border", align_corners=False)
    m = torch.sigmoid(hb[:, 2 + c:])
    rp = m * rw + (1 - m) * hb[:, 2:2 + c]
    assert rel_err(warped, rw) < 1e-6 and rel_err(pred, rp) < 1e-6
    w1, w2 = torch.randn_like(rw), torch.randn_like(rp)
    ((warped * w1).sum() + (pred * w2).sum()).backward()
    ((rw * w1).sum() + (rp * w2).sum()).backward()
    assert rel_err(pa.grad, pb.grad) < 1e-5 and rel_err(ha.grad, hb.grad) < 1e-5


@pytest.mark.parametrize("method", ["rk4", "dopri5"])
def test_chain_maxima_feed_the_scaled_weight_gradient(fake_backend, method):
    """At the fused-chain shape (16x16, 64 -> 64) the chain launches leave max |a_l| / max |d_l| in ConvNetOp's maxima buffer and
    the weight gradients take them (odex_conv2d_wgrad_f16_scaled): the fake backend asserts that every pair it is handed IS an
    upper bound of the operands of that launch (merged RK stages included), and the gradients still match the oracle."""
    from ode_rl_b200 import kernels as K
    f = make_ode_func(odex, C=64, n_units=64, n_layers=2, scale=2.0)
    solver = odex.DiffEqSolver(f, method, odeint_rtol=1e-3, odeint_atol=1e-4)
    _, tp = timesteps(3, 3)
    torch.manual_seed(4)
    z0 = torch.randn(2, 64, 16, 16, requires_grad=True)
    be = K.backend()
    sol = solver(z0, tp)
    w = torch.randn_like(sol)
    (sol * w).sum().backward()
    assert be.wgrad_scaled > 0, "the weight gradients did not take the chain's maxima"
    got = {n: p.grad.clone() for n, p in solver.named_parameters()}
    sd = _sd(solver)
    ps = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z1 = z0.detach().clone().requires_grad_(True)
    ref = refmodel.diffeq_solver(ps, z1, tp, method, 1e-3, 1e-4)
    (ref * w).sum().backward()
    assert rel_err(sol, ref) < 1e-5
    assert rel_err(z0.grad, z1.grad) < 1e-4
    for n, g in got.items():
        assert rel_err(g, ps[n].grad) < 1e-4, n


def test_chain_maxima_stay_tight_when_the_state_blows_up(fake_backend):
    """A field whose state grows by orders of magnitude over the rollout (and whose adjoint grows the other way): maxima shared
    over the whole backward would be far above the early steps' operands and flush their fp16 copies to zero.  The buffers are
    per RK step (rk4) / per evaluation (dopri5 backward), so every bound the weight gradients get is within 2^10 of the
    operand's own maximum (asserted by the fake backend) -- and two rollouts in flight do not disturb each other."""
    from ode_rl_b200 import kernels as K
    f = make_ode_func(odex, C=64, n_units=64, n_layers=2, scale=6.0)
    be = K.backend()
    for method in ("rk4", "dopri5"):
        solver = odex.DiffEqSolver(f, method, odeint_rtol=1e-3, odeint_atol=1e-4)
        _, tp = timesteps(3, 6)
        torch.manual_seed(5)
        za = torch.randn(1, 64, 16, 16, requires_grad=True)
        zb = (torch.randn(1, 64, 16, 16) * 30).requires_grad_(True)
        sa, sb = solver(za, tp), solver(zb, tp)
        assert float(sa[-1].abs().max()) > 1e3 * float(sa[0].abs().max())          # it does blow up
        n0 = be.wgrad_scaled
        sa.sum().backward()
        ga = {n: p.grad.clone() for n, p in solver.named_parameters()}
        for p in solver.parameters():
            p.grad = None
        sb.sum().backward()
        gb = {n: p.grad.clone() for n, p in solver.named_parameters()}
        assert be.wgrad_scaled > n0
        for z, got in ((za, ga), (zb, gb)):
            ps = {k: v.clone().requires_grad_(True) for k, v in _sd(solver).items()}
            ref = refmodel.diffeq_solver(ps, z.detach(), tp, method, 1e-3, 1e-4)
            ref.sum().backward()
            # rk4: same arithmetic on both sides.  dopri5 on a field this stiff: the two sides' step sizes differ in the last
            # digits and a few ReLU pre-activations sit on either side of 0 (tests/test_relu_replay.py) -- a sanity bound only;
            # what this test pins are the bound assertions inside the fake backend
            tol = 1e-4 if method == "rk4" else 2e-2
            for n, g in got.items():
                assert rel_err(g, ps[n].grad) < tol, (method, n)
        for p in solver.parameters():
            p.grad = None


def test_weight_gradient_launches_merge_over_stages_and_steps(fake_backend):
    """The arenas that make a layer's weight gradients consecutive slices (solver.RolloutArenas, the dopri5 backward's per-step
    arenas): rk4 on a 10-point grid = 36 evaluations of the chain, yet every layer but the first runs ONE launch over all 36 x B
    images and the first two (its stage-0 input is the solution buffer) -- and dopri5 one launch per layer and accepted step
    (+ the FSAL stage's own gradient plane for the last layer, + the first evaluation)."""
    from ode_rl_b200 import kernels as K
    be = K.backend()
    calls = []
    orig = be.conv2d_wgrad_packed

    def spy(x0, x1, dy, R, dwp, db, **kw):
        calls.append((x0.shape[0], kw.get("amax") is not None))
        return orig(x0, x1, dy, R, dwp, db, **kw)

    be.conv2d_wgrad_packed = spy
    try:
        f = make_ode_func(odex, C=64, n_units=64, n_layers=2)
        L = len(f.op().layers)
        _, tp = timesteps(10, 10)
        torch.manual_seed(0)
        z0 = torch.randn(2, 64, 16, 16, requires_grad=True)
        solver = odex.DiffEqSolver(f, "rk4")
        solver(z0, tp).sum().backward()
        assert sorted(calls) == sorted([(72, True)] * (L - 1) + [(54, True), (18, True)]), calls
        del calls[:]
        solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
        solver(z0, tp).sum().backward()
        n_acc = solver.last_stats.n_accept
        big = [c for c in calls if c[0] >= 10]
        assert len(big) == L * n_acc and all(c[1] for c in big), (n_acc, calls)     # 5-6 stages x B = 2 merged per layer and step
        assert len(calls) <= L * n_acc + 2 * n_acc + L, (n_acc, calls)
    finally:
        be.conv2d_wgrad_packed = orig
