import sys, torch
sys.path.insert(0, '.')
import ode_rl_b200 as odex
from tests.util import make_ode_func, timesteps
dev = torch.device('cuda')
f = make_ode_func(odex, C=64, n_units=64).to(dev)
cell = odex.ODEConvGRUCell(f, None, (16, 16), 64).to(dev)
obs, tp = timesteps(4, 4)
for method in ('euler', 'rk4', 'dopri5'):
    solver = odex.DiffEqSolver(f, method, odeint_rtol=1e-3, odeint_atol=1e-4)
    for B in (0, 1, 149):
        try:
            z0 = torch.randn(B, 64, 16, 16, device=dev, requires_grad=True)
            sol = solver(z0, tp.to(dev))
            sol.sum().backward()
            torch.cuda.synchronize()
            print(method, 'B=%d' % B, 'ok', tuple(sol.shape), bool(torch.isfinite(sol).all()))
        except Exception as e:
            print(method, 'B=%d' % B, 'FAIL', type(e).__name__, str(e)[:150])
    try:
        z0 = torch.randn(2, 64, 16, 16, device=dev)
        sol = solver(z0, tp[:1].to(dev))
        print(method, 'T=1 ok', tuple(sol.shape), float((sol[0] - z0).abs().max()))
    except Exception as e:
        print(method, 'T=1 FAIL', type(e).__name__, str(e)[:150])
for B in (0, 1):
    try:
        x = torch.randn(4, B, 64, 16, 16, device=dev, requires_grad=True)
        m, s = cell(x, obs.to(dev))
        (m.sum() + s.sum()).backward()
        torch.cuda.synchronize()
        print('cell B=%d ok' % B, tuple(m.shape))
    except Exception as e:
        print('cell B=%d FAIL' % B, type(e).__name__, str(e)[:150])
