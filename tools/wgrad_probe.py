"""Weight-gradient kernels at the shapes of the config-A step: 3xTF32 (odex_conv2d_wgrad_packed), kind::f16 with its own max
pass (odex_conv2d_wgrad_f16) and kind::f16 with caller-provided maxima (odex_conv2d_wgrad_f16_scaled); time per launch and
the relative difference to the 3xTF32 result.  ODEX_W16_TAPS3=n sets the taps per CTA of the 3x3 f16 kernel.
usage: python tools/wgrad_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()


def timed(fn, n=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


for (B, c0, c1, N, R) in [(512, 64, 0, 64, 3), (128, 64, 0, 64, 3), (128, 64, 64, 128, 5), (128, 64, 64, 64, 5)]:
    torch.manual_seed(0)
    x0 = torch.randn(B, 16, 16, c0, device="cuda")
    x1 = torch.randn(B, 16, 16, c1, device="cuda") if c1 else None
    dy = torch.randn(B, 16, 16, N, device="cuda") * 1e-3
    pm = lambda t: t.abs().amax(dim=(1, 2, 3))      # noqa: E731
    amax = ((pm(x0) if x1 is None else torch.maximum(pm(x0), pm(x1))).contiguous(), pm(dy).contiguous())
    res = {}
    for name, kw in (("tf32", {}), ("f16", {"f16": True}), ("f16_scaled", {"amax": amax}), ("f16_scaled_x64", {"amax": (amax[0] * 64, amax[1] * 64)})):
        dwp = torch.zeros(R * R, N, c0 + c1, device="cuda")
        db = torch.zeros(N, device="cuda")
        be.conv2d_wgrad_packed(x0, x1, dy, R, dwp, db, **kw)
        res[name] = (dwp.clone(), timed(lambda: be.conv2d_wgrad_packed(x0, x1, dy, R, dwp, db, **kw)))
    ref = res["tf32"][0]
    print("B=%d R=%d C=%d N=%d taps3=%s: " % (B, R, c0 + c1, N, os.environ.get("ODEX_W16_TAPS3", "5")) + "  ".join(
        "%s %.1f us (rel %.1e)" % (k, v[1], float((v[0] - ref).abs().max() / ref.abs().max())) for k, v in res.items()))
