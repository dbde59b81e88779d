import copy, sys, torch
sys.path.insert(0, '.')
from ode_rl_b200.harness import ODEConvGRUHarness
from tests.util import timesteps, rel_err
dev = torch.device("cuda")
torch.manual_seed(0)
m0 = ODEConvGRUHarness("rk4", device=dev).to(dev)
obs, tp = timesteps(4, 4, dev)
g = torch.Generator().manual_seed(7)
x = torch.rand(2, 4, 1, 64, 64, generator=g).to(dev)
y = torch.rand(2, 4, 1, 64, 64, generator=g).to(dev)
def run(opt_name):
    m = copy.deepcopy(m0)
    o = torch.optim.Adam(m.parameters(), lr=1e-3) if opt_name == "adam" else torch.optim.SGD(m.parameters(), lr=1e-1)
    for _ in range(5):
        o.zero_grad(set_to_none=True)
        loss = torch.nn.functional.mse_loss(m(x, obs, tp, return_latents=False), y)
        loss.backward()
        o.step()
    grads = {n: p.grad.clone() for n, p in m.named_parameters()}
    return m, grads
for opt_name in ("adam", "sgd"):
    a, ga = run(opt_name); b, gb = run(opt_name)
    worst = max((rel_err(p1, p2), n1) for (n1, p1), (_, p2) in zip(a.named_parameters(), b.named_parameters()))
    worstg = max((rel_err(ga[n], gb[n]), n) for n in ga)
    print(opt_name, "eager vs eager: worst param rel diff %.3e (%s); worst last-grad rel diff %.3e (%s)" % (worst + worstg))
