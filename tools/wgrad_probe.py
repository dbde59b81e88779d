"""Where does a wgrad_tc tile go?  Times the batched 3x3 / 5x5 launches of the step with parts of the kernel switched off
(ODEX_WGRAD_PROBE bitmask, results invalid): run once per mask.  usage: ODEX_WGRAD_PROBE=m python tools/wgrad_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
for (B, c0, c1, N, R) in [(512, 64, 0, 64, 3), (128, 64, 64, 128, 5)]:
    x0 = torch.randn(B, 16, 16, c0, device="cuda")
    x1 = torch.randn(B, 16, 16, c1, device="cuda") if c1 else None
    dy = torch.randn(B, 16, 16, N, device="cuda")
    dwp = torch.zeros(R * R, N, c0 + c1, device="cuda")
    db = torch.zeros(N, device="cuda")
    for _ in range(3):
        be.conv2d_wgrad_packed(x0, x1, dy, R, dwp, db)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        be.conv2d_wgrad_packed(x0, x1, dy, R, dwp, db)
    e1.record()
    torch.cuda.synchronize()
    print("probe=%s B=%d R=%d N=%d: %.1f us per launch" % (os.environ.get("ODEX_WGRAD_PROBE", "0"), B, R, N, e0.elapsed_time(e1) * 1e3 / 30))
