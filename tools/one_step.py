"""One eager training step of the bench workload bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off ...` (launch lists and --set full captures of exactly one step).
usage: python tools/one_step.py [method] [batch]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ode_rl_b200.harness import ODEConvGRUHarness  # noqa: E402


def main():
    method = sys.argv[1] if len(sys.argv) > 1 else "rk4"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = ODEConvGRUHarness(method, rtol=1e-3, atol=1e-4, device=dev).to(dev)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    x, y = bench.synth(B, 1234, dev)
    obs, tp = bench.times(dev)

    def step():
        opt.zero_grad(set_to_none=True)
        pred = model(x, obs, tp, return_latents=False)
        loss = torch.nn.functional.mse_loss(pred, y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    loss = step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("loss %.6f" % float(loss))


if __name__ == "__main__":
    main()
