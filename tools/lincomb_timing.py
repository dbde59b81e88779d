"""K3 (odex_lincomb) at the plane size of the config-A training step (B = 128: 8.4 MB per plane): us per launch and GB/s for
1 .. 6 operands, launches back to back with rotating buffers (no L2 reuse).  python tools/lincomb_timing.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ode_rl_b200 import kernels as K
be = K.Backend()
for B in (128, 512):
    n = B * 16 * 16 * 64
    pool = [torch.randn(n, device="cuda") for _ in range(40)]          # 40 x 8.4 MB = 336 MB > L2
    for nin in (1, 2, 3, 4, 5, 6):
        def run(k):
            ins = [pool[(k * 7 + j) % 40] for j in range(nin)]
            be.lincomb(pool[(k * 7 + nin) % 40], ins, [0.5] * nin)
        for k in range(5):
            run(k)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()                           # replayed: the host's ~12 us per ctypes launch must not be what is timed
        with torch.cuda.graph(g):
            for k in range(100):
                run(k)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 10
        print("B=%d nin=%d: %.2f us per launch, %.0f GB/s" % (B, nin, us, (nin + 1) * n * 4 / us / 1e3))
