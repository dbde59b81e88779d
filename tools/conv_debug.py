"""Role-level cycle counters of conv_tc2_kernel (CTA 0) + event-timed launches of v1 / v2 at the bench shape.
ODEX_CONV_DEBUG=1 python tools/conv_debug.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
for (B, c0, N, R) in [(128, 64, 64, 3), (512, 64, 64, 3), (128, 64, 64, 1)]:
    x0 = torch.randn(B, 16, 16, c0, device="cuda")
    w = torch.randn(N, c0, R, R, device="cuda") * 0.05
    bias = torch.randn(N, device="cuda")
    out = torch.empty(B, 16, 16, N, device="cuda")
    wt = be.pack_weights(w, False, True)
    for _ in range(3):
        be.conv2d(x0, None, None, wt, N, R, bias, K.ACT_RELU, out, impl=K.CONV_TC)
    torch.cuda.synchronize()
    if not os.environ.get("ODEX_CONV_DEBUG"):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            be.conv2d(x0, None, None, wt, N, R, bias, K.ACT_RELU, out, impl=K.CONV_TC)
        e1.record()
        torch.cuda.synchronize()
        print("B=%d R=%d: %.2f us per launch (back to back)" % (B, R, e0.elapsed_time(e1) * 1e3 / 50))
