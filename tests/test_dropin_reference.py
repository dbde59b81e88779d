"""The drop-in claim of SURVEY.md 8(b), tested literally: the reference's OWN caller files models/ODEConvGRU.py and
models/VidODE.py are executed UNMODIFIED with `modules.DiffEqSolver`, `modules.ODEConvGRUCell` and `modules.ConvGRUCell`
resolved to ode_rl_b200.modules (a sys.modules shim -- what a maintainer gets by putting the package on the path under
those names), and compared with the same files running over the reference's own modules:

  * same constructor call, identical state_dict keys; the reference state_dict loads strict=True;
  * forward (euler / rk4 / dopri5) equals the committed golden fixtures and the live reference;
  * backward reaches every parameter and the gradients agree with autograd through the reference modules.

Runs where /root/reference is mounted (the build container), on the CPU through tests/fake_backend.py (host logic + module
surface); the GPU kernels underneath are covered by tests/test_gpu_parity.py.  Skipped on the GPU box.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import pytest
import torch

from oracle import refload
from tests.test_golden import load_case
from tests.util import rel_err, rel_l2, timesteps

pytestmark = pytest.mark.skipif(not refload.available(), reason="reference tree not mounted")

_SHIMMED = ("modules.DiffEqSolver", "modules.ODEConvGRUCell", "modules.ConvGRUCell")


def _load_model_file_over_product(relpath: str, name: str):
    """Execute /root/reference/<relpath> as module `name` with the three hot-path module names bound to the product."""
    import ode_rl_b200.modules as prod

    refload.load()                                  # helpers.utils etc. importable, torchdiffeq / skimage stubbed
    saved = {k: sys.modules.get(k) for k in _SHIMMED}
    try:
        for k in _SHIMMED:
            m = types.ModuleType(k)
            for attr in ("DiffEqSolver", "ODEFunc", "ODEConvGRUCell", "ConvGRUCell"):
                setattr(m, attr, getattr(prod, attr))
            sys.modules[k] = m
        spec = importlib.util.spec_from_file_location(name, os.path.join(refload.REFERENCE_ROOT, relpath))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod


@pytest.fixture(scope="module")
def dropin_odeconvgru():
    return _load_model_file_over_product("models/ODEConvGRU.py", "_dropin_models_ODEConvGRU").ODEConvGRU


@pytest.fixture(scope="module")
def dropin_vidode():
    return _load_model_file_over_product("models/VidODE.py", "_dropin_models_VidODE").VidODE


def test_reference_model_file_builds_over_product_modules(reference, dropin_odeconvgru):
    import ode_rl_b200.modules as prod

    torch.manual_seed(0)
    ours = dropin_odeconvgru(refload.default_opt("euler"), torch.device("cpu"))
    torch.manual_seed(0)
    ref = reference.ODEConvGRU(refload.default_opt("euler"), torch.device("cpu"))
    assert isinstance(ours.diffeq_solver, prod.DiffEqSolver) and isinstance(ours.ode_convgru_cell, prod.ODEConvGRUCell)
    assert isinstance(ours.ode_convgru_cell.cgru_cell, prod.ConvGRUCell) and isinstance(ours.ode_decoder_func, prod.ODEFunc)
    assert not isinstance(ref.diffeq_solver, prod.DiffEqSolver)
    a, b = ref.state_dict(), ours.state_dict()
    assert list(a.keys()) == list(b.keys()) and len(a) == 60
    for k in a:                                      # same constructor order -> same RNG stream -> same initial weights
        assert torch.equal(a[k], b[k]), k
    ours.load_state_dict(a, strict=True)


@pytest.mark.parametrize("name", ["euler", "rk4", "dopri5", "dopri5_x4"])
def test_reference_odeconvgru_forward_over_product_matches_golden(fake_backend, dropin_odeconvgru, name):
    rec, sd, frames = load_case(name)
    model = dropin_odeconvgru(refload.default_opt(rec["method"]), torch.device("cpu"))
    model.load_state_dict(sd, strict=True)
    model.eval()
    model.diffeq_solver.odeint_rtol, model.diffeq_solver.odeint_atol = rec["rtol"], rec["atol"]
    obs, tp = timesteps(rec["T"], rec["T"])
    with torch.no_grad():
        pred = model.get_prediction(frames, {"observed_tp": obs, "tp_to_predict": tp})
    assert pred.shape == rec["pred"].shape
    tol = 1e-5 if rec["method"] != "dopri5" else rec["rtol"]
    assert rel_err(pred, rec["pred"]) < tol
    if rec["method"] == "dopri5":
        st = model.diffeq_solver.last_stats
        assert [bool(a[3]) for a in st.attempts] == [a[3] for a in rec["attempts"]]
        assert st.n_feval == rec["n_feval"]


class _HookedReLU(torch.nn.Module):
    """Stands in for the nn.ReLU instances of the reference's conv stacks: same forward value, sub-gradient mask replayed
    from the product run (oracle.refmodel.ReluReplay; tests/test_relu_replay.py explains why)."""

    def __init__(self, holder):
        super().__init__()
        self.holder = holder

    def forward(self, x):
        return self.holder["hook"].relu(x)


def _hook_reference_relus(model, holder):
    """Every conv stack of the reference model that contains nn.ReLU (ODEFunc.gradient_net, transform_z0): announce each
    evaluation to the hook and route its ReLUs through it.  Shared stacks (ode_decoder_func is diffeq_solver.ode_func) once."""
    seen = set()
    for seq in [m for m in model.modules() if isinstance(m, torch.nn.Sequential)]:
        if id(seq) in seen or not any(isinstance(c, torch.nn.ReLU) for c in seq):
            continue
        seen.add(id(seq))
        for i, c in enumerate(seq):
            if isinstance(c, torch.nn.ReLU):
                seq[i] = _HookedReLU(holder)
        seq.register_forward_pre_hook(lambda mod, args: holder["hook"].begin())


@pytest.mark.parametrize("method", ["euler", "rk4", "dopri5"])
def test_reference_odeconvgru_backward_over_product_matches_reference(fake_backend, reference, dropin_odeconvgru, method):
    """Forward + backward of the unmodified models/ODEConvGRU.py: 40/40 parameters get a gradient and it equals autograd
    through the reference's own modules (+ the restated odeint) -- at 1e-4 once both sides use the same ReLU sub-gradient at
    pre-activations that are 0 to rounding (seed 3 has two of them: without the replay one layer is off by 1e-2)."""
    from oracle import refmodel
    from tests.util import dopri5_backward_index_map, record_relu_masks

    B, T = 2, 4
    torch.manual_seed(3)
    ref = reference.ODEConvGRU(refload.default_opt(method), torch.device("cpu"))
    ours = dropin_odeconvgru(refload.default_opt(method), torch.device("cpu"))
    ours.load_state_dict(ref.state_dict(), strict=True)
    frames = torch.rand(B, T, 1, 64, 64, generator=torch.Generator().manual_seed(7))
    target = torch.rand(B, T, 1, 64, 64, generator=torch.Generator().manual_seed(8))
    ts = torch.tensor(np.arange(0, 2 * T) / (2 * T))
    batch = {"observed_tp": ts[:T], "tp_to_predict": ts[T:]}
    with record_relu_masks() as rec:
        pred = ours.get_prediction(frames, batch)
        torch.nn.functional.mse_loss(pred, target).backward()       # get_loss() hard-codes .cuda() (models/ODEConvGRU.py:96)
    index_map = None
    if method == "dopri5":      # encoder (T evaluations) + transform_z0 were recorded in the forward, the solver's in its backward
        accepted = [bool(a[3]) for a in ours.diffeq_solver.last_stats.attempts]
        index_map = list(range(T + 1)) + dopri5_backward_index_map(accepted, n_before=T + 1)
    hook = refmodel.ReluReplay(rec.masks(index_map))
    _hook_reference_relus(ref, {"hook": hook})
    rpred = ref.get_prediction(frames, batch)
    torch.nn.functional.mse_loss(rpred, target).backward()
    assert rel_err(pred, rpred) < (1e-5 if method != "dopri5" else 1e-4)
    assert hook.n_elems > 5e5 and hook.n_flips <= 8 and hook.max_flip_rel < 1e-5, (hook.n_flips, hook.max_flip_rel)
    ref_params, our_params = dict(ref.named_parameters()), dict(ours.named_parameters())
    assert ref_params.keys() == our_params.keys() and len(our_params) == 40
    for k, p in our_params.items():
        assert p.grad is not None, k
        g = ref_params[k].grad
        assert rel_l2(p.grad, g) < 1e-4, (k, rel_l2(p.grad, g))


def test_reference_vidode_file_over_product(fake_backend, reference, dropin_vidode):
    """models/VidODE.py unmodified.  Its forward hands a batch-first tensor to the time-first cell (models/VidODE.py:101-105,
    SURVEY.md App. B), so it only runs when B == T: use that case.  Same seed -> same state_dict; forward and backward equal
    the reference's own modules under the same file."""
    B = T = 3
    S, C = 64, 3
    opt = refload.vidode_opt("dopri5", S, C)
    torch.manual_seed(0)
    ref = reference.VidODE(opt, torch.device("cpu"))
    torch.manual_seed(0)
    ours = dropin_vidode(opt, torch.device("cpu"))
    a, b = ref.state_dict(), ours.state_dict()
    assert list(a.keys()) == list(b.keys())
    for k in a:
        assert torch.equal(a[k], b[k]), k
    frames = torch.rand(B, T, C, S, S, generator=torch.Generator().manual_seed(11))
    ts = torch.tensor(np.arange(0, 2 * T) / (2 * T))
    batch = {"observed_tp": ts[:T], "tp_to_predict": ts[T:], "observed_mask": torch.ones(B, T, 1),
             "mask_predicted_data": torch.ones(B, T, 1)}
    from oracle import refmodel
    from tests.util import dopri5_backward_index_map, record_relu_masks

    def run(m):
        pred, extra = m(frames, batch)
        assert set(extra) == {"optical_flow", "warped_pred_x", "pred_intermediates", "pred_masks"}
        pred.square().mean().backward()
        return pred.detach()

    with record_relu_masks() as rec:
        ours_pred = run(ours)
    st = ours.diffeq_solver.last_stats
    assert st is not None and st.n_accept >= 1
    accepted = [bool(a[3]) for a in st.attempts]
    hook = refmodel.ReluReplay(rec.masks(list(range(T + 1)) + dopri5_backward_index_map(accepted, n_before=T + 1)))
    for f in (ref.ode_encoder_func, ref.ode_decoder_func, ref.encoder_z0):     # the BatchNorm + ReLU stacks are torch on both sides
        _hook_reference_relus(f, {"hook": hook})
    ref_pred = run(ref)
    assert ref_pred.shape == (B, T, C, S, S)
    assert rel_err(ours_pred, ref_pred) < 1e-4
    assert hook.n_flips <= 8 and hook.max_flip_rel < 1e-5
    rp, op = dict(ref.named_parameters()), dict(ours.named_parameters())
    missing = [k for k, p in op.items() if (p.grad is None) != (rp[k].grad is None)]
    assert not missing, missing
    checked = 0
    for k, p in op.items():
        if p.grad is None:
            continue
        if k.endswith(".bias") and k[:-4] + "weight" in op and op[k[:-4] + "weight"].grad is not None:
            # a bias added right before a training-mode BatchNorm has an exactly-zero gradient: rounding noise on both sides
            wn = float(op[k[:-4] + "weight"].grad.norm())
            if float(p.grad.norm()) < 1e-4 * wn and float(rp[k].grad.norm()) < 1e-4 * wn:
                continue
        assert rel_l2(p.grad, rp[k].grad) < 5e-4, (k, rel_l2(p.grad, rp[k].grad))
        checked += 1
    assert checked > 40
