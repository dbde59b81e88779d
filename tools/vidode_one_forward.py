"""One VidODE forward (BASELINE configs[4]: 128x128x3, dopri5, batch 32) bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off ...` (launch list and --set full captures of exactly one forward).
usage: python tools/vidode_one_forward.py [batch]"""
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
        for _ in range(2):
            model(x, batch)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        pred, _ = model(x, batch)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    print("mean %.6f" % float(pred.mean()), model.diffeq_solver.last_stats)


if __name__ == "__main__":
    main()
