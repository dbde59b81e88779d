"""Role-level cycle counters of the weight-gradient kernels (CTA 0) at the bench shapes: wgrad_tc_kernel (3xTF32) and
wgrad_f16_kernel (kind::f16, per-image maxima given).  ODEX_WGRAD_DEBUG=1 python tools/wgrad_debug.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
pm = lambda t: t.abs().amax(dim=(1, 2, 3))      # noqa: E731
for (B, c0, c1, N, R) in [(128, 64, 0, 64, 3), (128, 64, 64, 128, 5), (128, 64, 64, 64, 5), (512, 64, 0, 64, 3)]:
    x0 = torch.randn(B, 16, 16, c0, device="cuda")
    x1 = torch.randn(B, 16, 16, c1, device="cuda") if c1 else None
    dy = torch.randn(B, 16, 16, N, device="cuda")
    dwp = torch.zeros(R * R, N, c0 + c1, device="cuda")
    db = torch.zeros(N, device="cuda")
    amax = ((pm(x0) if x1 is None else torch.maximum(pm(x0), pm(x1))).contiguous(), pm(dy).contiguous())
    for kw in ({}, {"amax": amax}):
        for _ in range(3):
            be.conv2d_wgrad_packed(x0, x1, dy, R, dwp, db, **kw)
        torch.cuda.synchronize()
