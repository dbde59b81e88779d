"""Time the torch/cuDNN encoder and decoder layers of the harness one by one (fwd, bwd) at the bench shapes.
usage: python tools/encdec_timing.py [n_images]"""
import sys
import torch
import torch.nn as nn
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1280
dev = "cuda"
layers = [
    ("enc1 Conv2d(1,16,3,2,1)    64x64", nn.Conv2d(1, 16, 3, 2, 1), (n, 1, 64, 64), False),
    ("enc2 Conv2d(16,64,3,2,1)   32x32", nn.Conv2d(16, 64, 3, 2, 1), (n, 16, 32, 32), True),
    ("dec1 ConvT(64,32,4,2,1)    16x16", nn.ConvTranspose2d(64, 32, 4, 2, 1), (n, 64, 16, 16), True),
    ("dec2 ConvT(32,1,4,2,1)     32x32", nn.ConvTranspose2d(32, 1, 4, 2, 1), (n, 32, 32, 32), True),
]
for name, m, shp, need_dx in layers:
    m = m.to(dev)
    x = torch.randn(*shp, device=dev, requires_grad=need_dx)
    y = m(x)
    g = torch.randn_like(y)
    for _ in range(3):
        y = m(x); y.backward(g)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    reps = 10
    tf = tb = 0.0
    for _ in range(reps):
        e[0].record(); y = m(x); e[1].record(); y.backward(g); e[2].record()
        torch.cuda.synchronize()
        tf += e[0].elapsed_time(e[1]); tb += e[1].elapsed_time(e[2])
    print("%s  fwd %.3f ms  bwd %.3f ms" % (name, tf / reps, tb / reps))
