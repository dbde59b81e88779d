"""Per-kernel device-time breakdown of one VidODE forward (BASELINE configs[4]; CUPTI via torch.profiler).
Run on the B200: python tools/vidode_profile.py [batch]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ode_rl_b200.harness import VidODEHarness  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = VidODEHarness("dopri5", device=dev).to(dev)
    model.device = dev
    x = torch.rand(B, 10, 3, 128, 128, device=dev)
    obs, tp = bench.times(dev)
    batch = {"observed_tp": obs, "tp_to_predict": tp, "observed_mask": torch.ones(B, 10, 1, device=dev),
             "mask_predicted_data": torch.ones(B, 10, 1, device=dev)}
    with torch.no_grad():
        for _ in range(3):
            model(x, batch)
        torch.cuda.synchronize()
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            model(x, batch)
            torch.cuda.synchronize()
    rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
    rows.sort(key=lambda r: -r[2])
    tot = sum(r[2] for r in rows)
    print("device time total %.2f ms over %d kernels" % (tot / 1e3, sum(r[1] for r in rows)))
    for k, n, t in rows[:30]:
        print("%9.1f us %5.1f%% n=%5d avg=%7.1f  %s" % (t, 100 * t / tot, n, t / n, k[:100]))


if __name__ == "__main__":
    main()
