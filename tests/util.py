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


class record_relu_masks:
    """Context manager: records, for every conv-stack evaluation the product runs with save=True (ConvNetOp.fwd), the
    sub-gradient masks its backward will use (saved activation > 0), as logical-NCHW bool CPU tensors, in call order:
    self.evals[i] = [mask of ReLU 1, mask of ReLU 2, ...].  Feed to oracle.refmodel.ReluReplay."""

    def __init__(self):
        self.evals = []

    def __enter__(self):
        from ode_rl_b200 import ops

        self._ops = ops
        self._orig = ops.ConvNetOp.fwd
        rec = self

        def fwd(op, x, save=False, **kw):
            r = rec._orig(op, x, save=save, **kw)
            if save:
                saved = r[1]
                rec.evals.append([(s.detach() > 0).movedim(-1, -3).cpu() for s in saved[1:]])
            return r

        ops.ConvNetOp.fwd = fwd
        return self

    def __exit__(self, *a):
        self._ops.ConvNetOp.fwd = self._orig
        return False

    def masks(self, index_map=None):
        """{oracle evaluation index: masks}; index_map[i] = oracle index of the product's i-th recorded evaluation
        (identity when omitted: euler / rk4 / the encoder recurrence evaluate in the same order on both sides)."""
        if index_map is None:
            return dict(enumerate(self.evals))
        return {index_map[i]: m for i, m in enumerate(self.evals) if index_map[i] is not None}


def relu_replay_gradients(product_grads, oracle_run, recorder, index_map=None):
    """Three-way ReLU gradient comparison.  oracle_run() -> {name: grad} runs the CPU oracle forward +
    backward from fresh leaves.  Returns (plain, replay, hook): rel-L2 error of every product gradient against the oracle
    as is, against the oracle with the product's ReLU sub-gradient masks replayed, and the hook with the flip statistics."""
    from oracle import refmodel

    g_plain = oracle_run()
    hook = refmodel.ReluReplay(recorder.masks(index_map))
    prev = refmodel.set_relu_hook(hook)
    try:
        g_replay = oracle_run()
    finally:
        refmodel.set_relu_hook(prev)
    plain = {k: rel_l2(v, g_plain[k]) for k, v in product_grads.items()}
    replay = {k: rel_l2(v, g_replay[k]) for k, v in product_grads.items()}
    return plain, replay, hook


def param_aliases(model):
    """{named_parameters() name: every state_dict key that is the same tensor} (a shared submodule is registered twice)."""
    by_ptr = {}
    for k, v in model.state_dict(keep_vars=True).items():
        by_ptr.setdefault(v.data_ptr(), []).append(k)
    return {n: by_ptr[p.data_ptr()] for n, p in model.named_parameters()}


def dopri5_backward_index_map(accepted, n_before=0):
    """Oracle evaluation index of every conv-stack evaluation the product's dopri5 BACKWARD records (solver._Dopri5Fn.backward
    recomputes, for the accepted steps latest first, the 6 stage evaluations, and at the end k0 = f(y0)).  The oracle
    evaluates f0, the initial-step probe, then 6 stages per ATTEMPT (accepted or not), after `n_before` earlier evaluations."""
    idx = []
    acc = [i for i, a in enumerate(accepted) if a]
    for ai in reversed(acc):
        idx.extend(n_before + 2 + 6 * ai + s for s in range(6))
    idx.append(n_before)
    return idx
