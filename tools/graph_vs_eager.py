"""Graph-replayed vs eagerly enqueued training steps under SGD (linear in the gradients: differences stay at the level of
the kernels' own run-to-run noise) and Adam (normalises every element's update, so fp32 summation-order noise on
near-zero gradient elements is amplified to a fraction of lr).  usage: python tools/graph_vs_eager.py"""
import copy, sys, torch
sys.path.insert(0, '.')
from ode_rl_b200.harness import GraphedTrainStep, ODEConvGRUHarness
from tests.util import timesteps, rel_err
dev = torch.device("cuda")
obs, tp = timesteps(4, 4, dev)
g = torch.Generator().manual_seed(7)
x = torch.rand(2, 4, 1, 64, 64, generator=g).to(dev)
y = torch.rand(2, 4, 1, 64, 64, generator=g).to(dev)
for name in ("sgd", "adam"):
    torch.manual_seed(0)
    m1 = ODEConvGRUHarness("rk4", device=dev).to(dev)
    m2 = copy.deepcopy(m1)
    mk = (lambda m: torch.optim.SGD(m.parameters(), lr=0.05)) if name == "sgd" else (lambda m: torch.optim.Adam(m.parameters(), lr=1e-3, capturable=True))
    o1, o2 = mk(m1), mk(m2)
    step = GraphedTrainStep(m1, o1, torch.nn.functional.mse_loss, x, y, obs, tp, warmup=2)
    l1 = [float(step()) for _ in range(3)]
    l2 = []
    for _ in range(5):
        o2.zero_grad(set_to_none=True)
        loss = torch.nn.functional.mse_loss(m2(x, obs, tp, return_latents=False), y)
        loss.backward(); o2.step(); l2.append(float(loss))
    worst = max((rel_err(p1, p2), n1) for (n1, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()))
    moved = max(float((p1 - p0).abs().max() / p0.abs().max()) for p1, p0 in zip(m1.parameters(), copy.deepcopy(m2).parameters()))
    print(name, "losses", l1, l2[2:], "worst param rel diff %.3e (%s)" % worst)
