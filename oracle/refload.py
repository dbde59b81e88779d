"""TEST INFRASTRUCTURE ONLY -- imports the REAL reference modules from /root/reference (read-only,
present only in the build container) with the two missing packages stubbed:
  * skimage.metrics.structural_similarity  (imported at helpers/utils.py:7, unused on the path)
  * torchdiffeq.odeint                      -> oracle/torchdiffeq_021.py (restatement, parity unpinned)
Used by oracle/make_golden.py to generate tests/golden/*.pt and by tests that validate
oracle/refmodel.py against the real thing.  Never imported by the product; never used on the GPU box.
"""
from __future__ import annotations

import os
import sys
import types
from argparse import Namespace

REFERENCE_ROOT = os.environ.get("ODEX_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "modules", "DiffEqSolver.py"))


def load():
    """Returns a namespace with the reference classes (ODEConvGRU, ODEConvGRUCell, ConvGRUCell,
    DiffEqSolver, ODEFunc, create_convnet)."""
    if not available():
        raise RuntimeError("reference not mounted at %s" % REFERENCE_ROOT)
    from . import torchdiffeq_021

    sys.dont_write_bytecode = True
    if "skimage" not in sys.modules:
        sk = types.ModuleType("skimage")
        skm = types.ModuleType("skimage.metrics")
        skm.structural_similarity = lambda *a, **k: 0.0
        sk.metrics = skm
        sys.modules["skimage"] = sk
        sys.modules["skimage.metrics"] = skm
    if "torchdiffeq" not in sys.modules:
        td = types.ModuleType("torchdiffeq")
        td.odeint = torchdiffeq_021.odeint
        sys.modules["torchdiffeq"] = td
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib

    ns = Namespace()
    m = importlib.import_module("modules.DiffEqSolver")
    ns.DiffEqSolver, ns.ODEFunc = m.DiffEqSolver, m.ODEFunc
    ns.ODEConvGRUCell = importlib.import_module("modules.ODEConvGRUCell").ODEConvGRUCell
    ns.ConvGRUCell = importlib.import_module("modules.ConvGRUCell").ConvGRUCell
    ns.ODEConvGRU = importlib.import_module("models.ODEConvGRU").ODEConvGRU
    ns.create_convnet = importlib.import_module("helpers.utils").create_convnet
    ns.VidODE = importlib.import_module("models.VidODE").VidODE
    return ns


def default_opt(method="euler", resolution=64, in_channels=1):
    """configs.yaml:593-605 `train_mmnist_odecgru_len20_1ch` + defaults, as a Namespace."""
    return Namespace(resolution=resolution, n_downs=2, conv_encoder_out_ch=64, in_channels=in_channels,
                     n_ode_layers=3, neural_ode_n_units=64, neural_ode_decoder_out_ch=64,
                     decode_diff_method=method, mem=False, z_sample=False)


def vidode_opt(method="dopri5", resolution=128, in_channels=3):
    """The options models/VidODE.py:46-88 reads (n_layers=2 as in SURVEY.md 8: f = 128->64->64->64->128)."""
    return Namespace(resolution=resolution, n_downs=2, in_channels=in_channels, n_layers=2, decode_diff_method=method)
