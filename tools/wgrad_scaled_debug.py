"""Every weight-gradient launch of a dopri5 + ReLU backward (the shape of tests/test_gpu_parity.py::
test_dopri5_relu_gradients_with_rejected_steps) run three ways -- scaled kind::f16 with the chain's maxima, kind::f16 with its own
max pass, 3xTF32 -- with the bounds it was handed beside the true maxima.  usage: python tools/wgrad_scaled_debug.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ode_rl_b200 as odex
from ode_rl_b200 import kernels as K
from tests.util import make_ode_func, timesteps

dev = torch.device("cuda")
f = make_ode_func(odex, C=64, n_units=64, scale=5.0, nonlinear="relu").to(dev)
solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
_, tp = timesteps(10, 10)
torch.manual_seed(2)
z0 = torch.randn(4, 64, 16, 16).to(dev).requires_grad_(True)
be = K.backend()
orig = be.conv2d_wgrad_packed
n = [0]
worst = [0.0]


def spy(x0, x1, dy, R, dwp, db, impl=0, f16=False, amax=None):
    outs = {}
    for name, kw in (("tf32", {}), ("f16", {"f16": True}), ("scaled", {"amax": amax} if amax is not None else None)):
        if kw is None:
            continue
        g, b = torch.zeros_like(dwp), (torch.zeros_like(db) if db is not None else None)
        orig(x0, x1, dy, R, g, b, **kw)
        outs[name] = g
    ref = outs["tf32"]
    msg = "call %3d B=%d |x| %.3e |dy| %.3e |ref| %.3e xptr %x" % (n[0], x0.shape[0], float(x0.abs().max()), float(dy.abs().max()), float(ref.abs().max()), x0.data_ptr())
    if amax is not None:
        msg += " bounds %.3e %.3e" % (float(amax[0]), float(amax[1]))
    for k in ("f16", "scaled"):
        if k in outs:
            msg += "  %s rel %.1e" % (k, float((outs[k] - ref).abs().max() / ref.abs().max().clamp_min(1e-30)))
    worst[0] = max(worst[0], max(float((outs[k] - ref).abs().max() / ref.abs().max().clamp_min(1e-30)) for k in outs))
    bad = any(float((outs[k] - ref).abs().max() / ref.abs().max().clamp_min(1e-30)) > 1e-4 for k in outs)
    if n[0] < 4 or bad or (amax is not None and (float(amax[0]) < float(x0.abs().max()) or float(amax[1]) < float(dy.abs().max()))):
        print(msg)
    n[0] += 1
    return orig(x0, x1, dy, R, dwp, db, impl=impl, f16=f16, amax=amax)


def run(scaled, spied):
    os.environ["ODEX_WGRAD_SCALED"] = "1" if scaled else "0"
    be.conv2d_wgrad_packed = spy if spied else orig
    for p_ in solver.parameters():
        p_.grad = None
    torch.manual_seed(7)
    sol = solver(z0, tp.to(dev))
    (sol * torch.randn_like(sol)).sum().backward()
    torch.cuda.synchronize()
    return {k: v.grad.clone() for k, v in solver.named_parameters()}


g_spy = run(True, True)
print("calls", n[0], "worst per-launch rel", worst[0], solver.last_stats)
g_off = run(False, False)
g_on = run(True, False)
g_on2 = run(True, False)
for k in g_off:
    den = g_off[k].abs().max()
    print(k, "on vs off %.2e  spied vs off %.2e  on vs on %.2e" % (float((g_on[k] - g_off[k]).abs().max() / den),
          float((g_spy[k] - g_off[k]).abs().max() / den), float((g_on[k] - g_on2[k]).abs().max() / den)))
