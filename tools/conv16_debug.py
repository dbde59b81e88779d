"""odex_conv16 at the bench shape (B = 128, 16x16, 5x5): event-timed launches per mode, or CTA-0 cycle marks with ODEX_CONV_DEBUG=1.
usage: [ODEX_CONV_DEBUG=1] python tools/conv16_debug.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
C = 64
x, h = torch.randn(B, 16, 16, C, device="cuda"), torch.randn(B, 16, 16, C, device="cuda")
wg = be.pack_weights(torch.randn(2 * C, 2 * C, 5, 5, device="cuda") / 55.0, False, 2)
wc = be.pack_weights(torch.randn(C, 2 * C, 5, 5, device="cuda") / 55.0, False, 2)
wd1 = be.pack_weights(torch.randn(C, 2 * C, 5, 5, device="cuda") / 55.0, True, 2)       # 64 -> 128
wd2 = be.pack_weights(torch.randn(2 * C, 2 * C, 5, 5, device="cuda") / 55.0, True, 2)   # 128 -> 128
w3 = be.pack_weights(torch.randn(C, C, 3, 3, device="cuda") / 24.0, False, 2)
bg, gam, bet = torch.randn(2 * C, device="cuda") * 0.1, torch.rand(2 * C, device="cuda") + 0.5, torch.randn(2 * C, device="cuda") * 0.1
z, rh, st = torch.empty_like(x), torch.empty_like(x), torch.empty(B * 8, device="cuda")
g = torch.empty(B, 16, 16, 2 * C, device="cuda")
hn, cand, cpre = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
D1, D2 = torch.empty(B, 16, 16, 2 * C, device="cuda"), torch.empty(B, 16, 16, 2 * C, device="cuda")
o3 = torch.empty_like(x)
cases = {
    "gates (save)": lambda: be.conv16_gru_gates(x, h, wg, 5, bg, gam, bet, 1e-5, z, rh, st, g_raw=g),
    "gates (inference)": lambda: be.conv16_gru_gates(x, h, wg, 5, bg, gam, bet, 1e-5, z, rh, st),
    "out (save)": lambda: be.conv16_gru_out(x, rh, wc, 5, bg[:C], gam[:C], bet[:C], 1e-5, z, h, hn, st, cand=cand, cpre=cpre),
    "out (inference)": lambda: be.conv16_gru_out(x, rh, wc, 5, bg[:C], gam[:C], bet[:C], 1e-5, z, h, hn, st),
    "dgrad 64->128": lambda: be.conv16(x, None, wd1, 2 * C, 5, None, K.ACT_NONE, D1),
    "dgrad 128->128 + res": lambda: be.conv16(g[..., :C], g[..., C:], wd2, 2 * C, 5, None, K.ACT_NONE, D2, aux=D1, aux_kind=3),
    "3x3 64->64": lambda: be.conv16(x, None, w3, C, 3, bg[:C], K.ACT_RELU, o3),
}
for name, fn in cases.items():
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if os.environ.get("ODEX_CONV_DEBUG"):
        continue
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print("B=%d %-22s %.1f us per launch" % (B, name, e0.elapsed_time(e1) * 1e3 / 30))
