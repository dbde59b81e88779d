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


if __name__ == "__main__":
    main()
