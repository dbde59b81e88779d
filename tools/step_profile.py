"""Per-kernel device-time breakdown of one training step (CUPTI via torch.profiler; no ncu serialisation) plus the
host enqueue time, to tell a GPU-bound step from a launch-bound one.  Run on the B200: python tools/step_profile.py"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ode_rl_b200.harness import ODEConvGRUHarness  # noqa: E402


def main():
    method = sys.argv[1] if len(sys.argv) > 1 else "rk4"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cudnn.benchmark = os.environ.get('CUDNN_BENCHMARK', '0') == '1'
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

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("host enqueue %.2f ms, until device idle %.2f ms" % ((t1 - t0) * 1e3, (t2 - t0) * 1e3))
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
        torch.cuda.synchronize()
    rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
    rows.sort(key=lambda r: -r[2])
    tot = sum(r[2] for r in rows)
    print("device time total %.2f ms over %d kernels" % (tot / 1e3, sum(r[1] for r in rows)))
    for k, n, t in rows[:28]:
        print("%9.1f us %5.1f%% n=%5d avg=%7.1f  %s" % (t, 100 * t / tot, n, t / n, k[:90]))


if __name__ == "__main__":
    main()
