"""Shared test helpers: ABI-layout synthetic problems <-> oracle structures."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from meta_pde_b200 import synthetic  # noqa: E402
from meta_pde_b200._abi import PDE_OUT_DIM, PDE_TERMS  # noqa: E402


def oracle_points(pde, sets, t, dtype=torch.float64):
    """ABI point sets of task t -> the tuple the reference's loss_fn takes."""
    s = [x[t].to(dtype) if x.dim() == 3 else x.to(dtype) for x in sets]
    if pde == "poisson":
        return (s[0], s[1])
    if pde == "td_burgers":
        return (s[0], s[1], s[2], s[3])
    # reference tuple: top, bottom, left, right, holes, domain (left/right/holes unused by the loss)
    return (s[0], s[1], s[0], s[0], s[0], s[2])


def oracle_params(pde, prm, t, dtype=torch.float64):
    p = prm[t].to(dtype)
    if pde == "poisson":
        return (p[0:6].reshape(2, 3), p[6:11], p[11:13])
    if pde == "td_burgers":
        return (p[0:1], p[1:3])
    return (p[0:2], p[2:4], torch.zeros(25, 5, dtype=dtype), 25)


def rel_l2(a, b):
    a = a.double().reshape(-1)
    b = b.double().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-300))
