"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.pt by running the REAL reference modules
(models/ODEConvGRU.py, modules/*.py imported read-only from /root/reference) on CPU fp32 with seeded synthetic
Moving-MNIST-shaped inputs.  `torchdiffeq.odeint` is the restatement in oracle/torchdiffeq_021.py
(the real package is absent: parity of the solver arithmetic itself stays unpinned, see that file's header).
Run in the build container only:   python -m oracle.make_golden
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refload, torchdiffeq_021 as tde  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    ref = refload.load()
    os.makedirs(OUT, exist_ok=True)
    B, T = 2, 10
    torch.manual_seed(0)
    model = ref.ODEConvGRU(refload.default_opt("euler"), torch.device("cpu"))
    model.eval()
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    torch.save(sd, os.path.join(OUT, "odeconvgru_state_dict.pt"))
    frames = torch.rand(B, T, 1, 64, 64, generator=torch.Generator().manual_seed(1234))
    ts = torch.tensor(np.arange(0, 2 * T) / (2 * T))
    batch = {"observed_tp": ts[:T], "tp_to_predict": ts[T:]}
    for method, scale in (("euler", 1.0), ("rk4", 1.0), ("dopri5", 1.0), ("dopri5", 3.0), ("dopri5", 4.0)):
        model.load_state_dict(sd)
        if scale != 1.0:    # stiffer decoder ODE so that dopri5 rejects steps
            with torch.no_grad():
                for p in model.ode_decoder_func.parameters():
                    p.mul_(scale)
        model.diffeq_solver.ode_method = method
        model.diffeq_solver.odeint_rtol, model.diffeq_solver.odeint_atol = 1e-3, 1e-4
        with torch.no_grad():
            enc = model.conv_encoder(frames.view(B * T, 1, 64, 64)).view(B, T, 64, 16, 16).permute(1, 0, 2, 3, 4)
            mean_z0, std_z0 = model.ode_convgru_cell(enc, batch["observed_tp"])
            sol = model.diffeq_solver(mean_z0, batch["tp_to_predict"])
            log = tde.LAST_LOG
            pred = model(frames, batch)
        rec = {"method": method, "scale": scale, "rtol": 1e-3, "atol": 1e-4, "frames_seed": 1234, "B": B, "T": T,
               "mean_z0": mean_z0.contiguous(), "std_z0": std_z0.contiguous(),
               "sol_last": sol[-1].clone(), "sol_mid": sol[T // 2].clone(), "pred": pred.contiguous(),
               "attempts": list(log.attempts) if method == "dopri5" else [],
               "n_feval": log.n_feval if method == "dopri5" else None}
        name = "odeconvgru_%s%s.pt" % (method, "" if scale == 1.0 else "_x%d" % int(scale))
        torch.save(rec, os.path.join(OUT, name))
        print(name, "pred", tuple(pred.shape), "attempts", len(rec["attempts"]),
              "rejects", sum(1 for a in rec["attempts"] if not a[3]))
    # one ConvGRUCell step and one ODEFunc evaluation of the reference classes, with their own default init
    torch.manual_seed(1)
    cell = ref.ConvGRUCell((16, 16), 64, 64, 5)
    x, h = torch.randn(1, 2, 64, 16, 16), torch.randn(2, 64, 16, 16)
    with torch.no_grad():
        out, hn = cell(x, h, seq_len=1)
    torch.save({"state_dict": cell.state_dict(), "x": x, "h": h, "h_next": hn}, os.path.join(OUT, "convgrucell_step.pt"))
    f = ref.ODEFunc(64, 64, 3, 64, False, "relu", final_act=False)
    y = torch.randn(2, 64, 16, 16)
    with torch.no_grad():
        fy = f(0.0, y)
    torch.save({"state_dict": f.state_dict(), "y": y, "f": fy}, os.path.join(OUT, "odefunc_eval.pt"))


def vidode_forward(model, frames, batch):
    """models/VidODE.py:91-141 through the reference's OWN submodules and methods, with the two layout fixes of
    SURVEY.md Appendix B (time-first tensor into encoder_z0; time-first solution permuted to batch-first)."""
    b, t, c, h, w = frames.size()
    enc = model.conv_encoder(frames.view(b * t, c, h, w))
    enc = enc.view(b, t, *enc.shape[-3:]).permute(1, 0, 2, 3, 4)
    mu, _ = model.encoder_z0(enc, batch["observed_tp"], batch["observed_mask"])
    sol = model.diffeq_solver(mu, batch["tp_to_predict"])
    log = tde.LAST_LOG
    sol_y = sol.permute(1, 0, 2, 3, 4).contiguous()
    skip = model.conv_encoder(frames[:, -1, ...])
    outs = torch.cat(model.get_flowmaps(sol_out=sol_y, first_prev_embed=skip, mask=batch["mask_predicted_data"]), dim=1)
    nc = model.opt.in_channels
    flows, inter, masks = outs[:, :, :2, ...], outs[:, :, 2:2 + nc, ...], torch.sigmoid(outs[:, :, 2 + nc:, ...])
    gx = torch.linspace(-1.0, 1.0, w).view(1, 1, w, 1).expand(b, h, -1, -1)
    gy = torch.linspace(-1.0, 1.0, h).view(1, h, 1, 1).expand(b, -1, w, -1)
    grid = torch.cat([gx, gy], 3).float()
    warped = torch.cat(model.get_warped_images(pred_flows=flows, start_image=frames[:, -1, ...], grid=grid), dim=1)
    pred = (masks * warped + (1 - masks) * inter).view(b, -1, c, h, w)
    return mu, sol_y, pred, log


def vidode_batch(B, T, C, S, seed=1234):
    frames = torch.rand(B, T, C, S, S, generator=torch.Generator().manual_seed(seed))
    ts = torch.tensor(np.arange(0, 2 * T) / (2 * T))
    return frames, {"observed_tp": ts[:T], "tp_to_predict": ts[T:], "observed_mask": torch.ones(B, T, 1),
                    "mask_predicted_data": torch.ones(B, T, 1)}


def param_checksums(model):
    return {k: float(v.double().sum()) for k, v in model.state_dict().items() if v.dtype.is_floating_point}


def main_vidode_128():
    """BASELINE configs[4]'s own frame size: 128x128x3 frames -> 128 latent channels on 32x32 (the shapes at which the kernels take
    their 32x32 / 256-gate paths), dopri5 at the DiffEqSolver defaults, B=1, 2 in / 2 out (the CPU reference needs seconds)."""
    ref = refload.load()
    B, T, C, S = 1, 2, 3, 128
    frames, batch = vidode_batch(B, T, C, S, seed=4321)
    torch.manual_seed(0)
    model = ref.VidODE(refload.vidode_opt("dopri5", S, C), torch.device("cpu"))
    sums = param_checksums(model)
    with torch.no_grad():
        mu, sol_y, pred, log = vidode_forward(model, frames, batch)
    rec = {"method": "dopri5", "scale": 1.0, "rtol": model.diffeq_solver.odeint_rtol, "atol": model.diffeq_solver.odeint_atol,
           "frames_seed": 4321, "B": B, "T": T, "C": C, "S": S, "weights_seed": 0, "checksums": sums,
           "mean_z0": mu.contiguous(), "sol_y": sol_y, "pred": pred.contiguous(), "attempts": list(log.attempts), "n_feval": log.n_feval}
    torch.save(rec, os.path.join(OUT, "vidode_dopri5_128.pt"))
    print("vidode_dopri5_128.pt pred", tuple(pred.shape), "attempts", len(rec["attempts"]))


def main_vidode():
    """BASELINE configs[4] wiring at a fixture-sized shape: 64x64x3 frames (16x16 latent, 128 channels), B=2, 3 in / 3 out.
    Weights are NOT stored (13 MB): the fixture keeps per-tensor checksums of the seeded default init and the product
    harness re-creates them from the same seed (tests/test_golden.py checks state_dict equality against the reference
    class where /root/reference is mounted, and the checksums everywhere)."""
    ref = refload.load()
    B, T, C, S = 2, 3, 3, 64
    frames, batch = vidode_batch(B, T, C, S)
    for method, scale in (("rk4", 1.0), ("dopri5", 1.0), ("dopri5", 4.0)):
        torch.manual_seed(0)
        model = ref.VidODE(refload.vidode_opt(method, S, C), torch.device("cpu"))     # train mode: BatchNorm uses batch statistics
        if scale != 1.0:
            with torch.no_grad():
                for p in model.ode_decoder_func.parameters():
                    p.mul_(scale)
        sums = param_checksums(model)          # before the forward: BatchNorm running statistics still at their init
        with torch.no_grad():
            mu, sol_y, pred, log = vidode_forward(model, frames, batch)
        rec = {"method": method, "scale": scale, "rtol": model.diffeq_solver.odeint_rtol, "atol": model.diffeq_solver.odeint_atol,
               "frames_seed": 1234, "B": B, "T": T, "C": C, "S": S, "weights_seed": 0, "checksums": sums,
               "mean_z0": mu.contiguous(), "sol_y": sol_y, "pred": pred.contiguous(),
               "attempts": list(log.attempts) if method == "dopri5" else [],
               "n_feval": log.n_feval if method == "dopri5" else None}
        name = "vidode_%s%s.pt" % (method, "" if scale == 1.0 else "_x%d" % int(scale))
        torch.save(rec, os.path.join(OUT, name))
        print(name, "pred", tuple(pred.shape), "attempts", len(rec["attempts"]),
              "rejects", sum(1 for a in rec["attempts"] if not a[3]))


if __name__ == "__main__":
    if "vidode128" in sys.argv[1:]:
        main_vidode_128()
    elif "vidode" in sys.argv[1:]:
        main_vidode()
    else:
        main()
