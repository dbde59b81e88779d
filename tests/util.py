"""Shared helpers for the tests: seeded synthetic inputs and model builders."""
import numpy as np
import torch


def timesteps(n_in=10, n_out=10, device="cpu"):
    ts = torch.tensor(np.arange(0, n_in + n_out) / (n_in + n_out))
    return ts[:n_in].to(device), ts[n_in:].to(device)


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max |a-b| / max |b| -- the 'relative per predicted latent' measure used for the 1e-5 bar."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def make_ode_func(ode_rl, C=64, n_layers=3, n_units=64, seed=0, scale=1.0, final_act=False, nonlinear="relu"):
    torch.manual_seed(seed)
    f = ode_rl.ODEFunc(n_inputs=C, n_outputs=C, n_layers=n_layers, n_units=n_units, downsize=False,
                       nonlinear=nonlinear, final_act=final_act)
    if scale != 1.0:
        with torch.no_grad():
            for p in f.parameters():
                p.mul_(scale)
    return f


def latent_rel(a: torch.Tensor, b: torch.Tensor) -> float:
    """north_star's bar, 'relative per predicted latent': worst over latents (leading dims = time/batch, last 3 =
    C,H,W) of ||a_i - b_i||_2 / ||b_i||_2."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    a, b = a.reshape(-1, *a.shape[-3:]).flatten(1), b.reshape(-1, *b.shape[-3:]).flatten(1)
    return float(((a - b).norm(dim=1) / b.norm(dim=1).clamp_min(1e-30)).max())
