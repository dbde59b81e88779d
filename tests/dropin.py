"""Builders shared by the parity tests."""
import torch

from ode_rl_b200.harness import ODEConvGRUHarness


def build_product_model(method, rtol, atol, state_dict, device):
    m = ODEConvGRUHarness(method, rtol, atol, device=device)
    m.load_state_dict(state_dict, strict=True)      # reference key names, aliased entries included
    return m.to(device).eval()
