"""Which torch ops are left in the training step, and who calls them: CPU-side aten ops with their device time and the Python
frames that issued them (torch.profiler with_stack).  Run on the B200: python tools/torch_side_hunt.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ode_rl_b200.harness import ODEConvGRUHarness  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = ODEConvGRUHarness("rk4", rtol=1e-3, atol=1e-4, device=dev).to(dev)
opt = torch.optim.Adam(model.parameters(), lr=1e-3)
x, y = bench.synth(128, 1234, dev)
obs, tp = bench.times(dev)


def step():
    opt.zero_grad(set_to_none=True)
    loss = torch.nn.functional.mse_loss(model(x, obs, tp, return_latents=False), y)
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages(group_by_input_shape=True):
    if e.device_time_total > 15 and e.key.startswith("aten::") and e.key not in ("aten::to", "aten::_to_copy", "aten::contiguous"):
        rows.append((e.device_time_total, e.count, e.key, str(e.input_shapes)[:150]))
rows.sort(reverse=True)
for t, n, k, st in rows[:40]:
    print("%8.1f us n=%3d %-28s %s" % (t, n, k, st))
