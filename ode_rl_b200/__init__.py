"""ode_rl_b200 -- B200-native (sm_100a) implementation of ODE-RL's latent Neural-ODE rollout hot path.

Exports the reference's module surface for that path (modules/DiffEqSolver.py, modules/ODEConvGRUCell.py,
modules/ConvGRUCell.py, helpers/utils.py:create_convnet); all compute runs in libodex.so (csrc/, C-ABI in
include/odex.h).  There is no CPU or eager-PyTorch fallback.
"""
from .modules import ConvGRUCell, DiffEqSolver, ODEConvGRUCell, ODEFunc, create_convnet  # noqa: F401
from .ops import set_conv_impl, set_conv_split  # noqa: F401
from ._lib import OdexError  # noqa: F401

__all__ = ["DiffEqSolver", "ODEFunc", "ODEConvGRUCell", "ConvGRUCell", "create_convnet", "set_conv_impl", "set_conv_split", "OdexError"]
