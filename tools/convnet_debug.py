"""Fused conv chain (odex_convnet) at the bench shape: CTA-0 cycle count (ODEX_CONV_DEBUG=1) or event-timed launches.
usage: [ODEX_CONV_DEBUG=1] python tools/convnet_debug.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
SPLIT = 0 if os.environ.get('ODEX_SPLIT', 'f16') == 'tf32' else 1
for (B, L, save) in [(128, 5, False), (128, 5, True), (512, 5, True)]:
    x = torch.randn(B, 16, 16, 64, device="cuda")
    ws = [be.pack_weights(torch.randn(64, 64, 3, 3, device="cuda") / 24.0, False, 1 + SPLIT) for _ in range(L)]
    bs = [torch.randn(64, device="cuda") for _ in range(L)]
    outs = [torch.empty_like(x) if (save or l == L - 1) else None for l in range(L)]
    acts = [K.ACT_RELU] * (L - 1) + [K.ACT_NONE]
    aux = [None] * (L - 1) + [x]
    kinds = [0] * (L - 1) + [3]
    for _ in range(3):
        be.convnet(x, ws, bs, acts, aux, kinds, outs, alpha=0.05, split=SPLIT)
    torch.cuda.synchronize()
    if not os.environ.get("ODEX_CONV_DEBUG"):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            be.convnet(x, ws, bs, acts, aux, kinds, outs, alpha=0.05, split=SPLIT)
        e1.record()
        torch.cuda.synchronize()
        print("B=%d L=%d save=%s: %.2f us per launch (%.2f per layer)" % (B, L, save, e0.elapsed_time(e1) * 1e3 / 50, e0.elapsed_time(e1) * 1e3 / 50 / L))
