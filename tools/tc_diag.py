"""GPU diagnostic for the tcgen05 conv kernel: each case in its own subprocess (a trap poisons the CUDA context),
compared with the exact-fp32 FFMA kernel and a torch fp64 CPU reference; prints where the error lives."""
import subprocess
import sys

CASES = ["2,16,16,64,0,64,3", "2,16,16,64,0,64,1", "2,16,16,64,0,128,3", "2,16,16,64,64,128,5", "2,16,16,64,64,64,5",
         "1,32,32,128,0,64,3", "1,8,8,64,0,64,3", "128,16,16,64,0,64,3"]


def run_case(spec):
    import torch
    import torch.nn.functional as F
    sys.path.insert(0, ".")
    from ode_rl_b200 import kernels as K
    B, H, W, c0, c1, N, R = [int(v) for v in spec.split(",")]
    be = K.Backend()
    torch.manual_seed(0)
    x0 = torch.randn(B, H, W, c0, device="cuda")
    x1 = torch.randn(B, H, W, c1, device="cuda") if c1 else None
    w = torch.randn(N, c0 + c1, R, R, device="cuda") / (R * (c0 + c1) ** 0.5)
    b = torch.randn(N, device="cuda")
    x = x0 if x1 is None else torch.cat((x0, x1), 3)
    ref = F.conv2d(x.permute(0, 3, 1, 2).double().cpu(), w.double().cpu(), b.double().cpu(), padding=(R - 1) // 2).permute(0, 2, 3, 1)
    outs = {}
    for name, tc in (("simt", False), ("tc", True)):
        if not tc and (W % 16 or H % 8):
            continue
        o = torch.full((B, H, W, N), float("nan"), device="cuda")
        wp = be.pack_weights(w, False, tc)
        be.conv2d(x0, x1, None if tc else wp, wp if tc else None, N, R, b, 0, o, impl=2 if tc else 1)
        torch.cuda.synchronize()
        outs[name] = o.double().cpu()
        e = (outs[name] - ref).abs()
        print("  %-4s max_abs_err %.3e  rel %.3e  nan %d" % (name, float(e.nan_to_num(1e9).max()), float(e.nan_to_num(1e9).max() / ref.abs().max()),
                                                         int(torch.isnan(outs[name]).sum())))
        if float(e.nan_to_num(1e9).max() / ref.abs().max()) > 1e-4:
            print("     err by channel block (32):", [round(float(e[..., i:i + 32].nan_to_num(1e9).mean()), 4) for i in range(0, N, 32)])
            print("     err by image row:", [round(float(e[0, r].nan_to_num(1e9).mean()), 4) for r in range(H)])
            print("     err by image col:", [round(float(e[0, :, c].nan_to_num(1e9).mean()), 4) for c in range(W)])
            print("     out[0,0,0,:4] %s ref %s" % (outs[name][0, 0, 0, :4].tolist(), ref[0, 0, 0, :4].tolist()))
            print("     out[0,5,7,:4] %s ref %s" % (outs[name][0, 5 % H, 7 % W, :4].tolist(), ref[0, 5 % H, 7 % W, :4].tolist()))
            ratio = (outs[name] / ref)
            print("     median out/ref %.4f" % float(ratio.nan_to_num(0).median()))
    # timing
    if "tc" in outs:
        for name, tc in (("simt", False), ("tc", True)):
            if not tc and (W % 16 or H % 8):
                continue
            wp = be.pack_weights(w, False, tc)
            o = torch.empty((B, H, W, N), device="cuda")
            for _ in range(3):
                be.conv2d(x0, x1, None if tc else wp, wp if tc else None, N, R, b, 0, o, impl=2 if tc else 1)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(20):
                be.conv2d(x0, x1, None if tc else wp, wp if tc else None, N, R, b, 0, o, impl=2 if tc else 1)
            e.record()
            torch.cuda.synchronize()
            us = s.elapsed_time(e) / 20 * 1e3
            fl = 2.0 * B * H * W * N * (c0 + c1) * R * R
            print("  %-4s %.1f us/launch  %.1f TFLOP/s (algorithmic)" % (name, us, fl / us / 1e6))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run_case(sys.argv[1])
    else:
        for c in CASES:
            print("case B,H,W,c0,c1,N,R =", c, flush=True)
            r = subprocess.run([sys.executable, __file__, c], capture_output=True, text=True, timeout=300)
            print(r.stdout, end="")
            if r.returncode != 0:
                print("  EXIT", r.returncode, r.stderr[-1500:])
