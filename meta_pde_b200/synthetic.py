"""Deterministic synthetic inputs for tests and bench.py.

Everything is produced on the CPU with seeded ``torch.Generator``s in the C-ABI
layout (flat theta, point sets ``[T, n, 2]``, ``pde_params [T, nprm]``); the
same tensors are handed to the CUDA path and, by tests/ and the bench's CPU
baseline leg, to the oracle.
"""
import math

import torch

from ._abi import PDE_NPRM, PDE_OUT_DIM, PDE_SETS


def siren_theta(sizes, out_dim=1, omega=30.0, omega0=30.0, log_scale=True, io_scale_lr_factor=10.0, seed=0,
                bias_scale=0.0, in_dim=2, dtype=torch.float32):
    """Flat theta with the reference's init distributions (src/nets/field.py:23-44,140-143):
    first layer (omega0/omega) U(+-1/fan_in), others U(+-sqrt(6/fan_in)/omega), zero biases
    (``bias_scale`` > 0 perturbs them for tests), log io-scales = log(1/io_scale_lr_factor)."""
    g = torch.Generator().manual_seed(int(seed))
    parts = []
    n_in = in_dim
    dims = list(sizes) + [out_dim]
    for l, n_out in enumerate(dims):
        u = torch.rand(n_in, n_out, generator=g, dtype=torch.float64) * 2.0 - 1.0
        if l == 0:
            k = (omega0 / omega) * u / n_in
        else:
            k = u * math.sqrt(6.0 / n_in) / omega
        parts.append(k.reshape(-1))
        parts.append(bias_scale * torch.randn(n_out, generator=g, dtype=torch.float64))
        n_in = n_out
    if log_scale:
        parts.append(torch.full((in_dim,), math.log(1.0 / io_scale_lr_factor), dtype=torch.float64))
        parts.append(torch.full((out_dim,), math.log(1.0 / io_scale_lr_factor), dtype=torch.float64))
    return torch.cat(parts).to(dtype)


def swish_theta(sizes, out_dim=1, log_scale=True, io_scale_lr_factor=10.0, seed=0, bias_scale=0.0, in_dim=2,
                dtype=torch.float32):
    """Flat theta of the `--nosiren` network (src/nets/field.py:171-176: variance_scaling(1.0, "fan_in",
    "truncated_normal") on every kernel): N(0, 1 / fan_in) truncated at 2 sigma, rescaled by 1 / 0.8796..."""
    g = torch.Generator().manual_seed(int(seed))
    parts = []
    n_in = in_dim
    for n_out in list(sizes) + [out_dim]:
        z = torch.randn(n_in, n_out, generator=g, dtype=torch.float64).clamp(-2.0, 2.0)
        parts.append((z * math.sqrt(1.0 / n_in) / 0.87962566103423978).reshape(-1))
        parts.append(bias_scale * torch.randn(n_out, generator=g, dtype=torch.float64))
        n_in = n_out
    if log_scale:
        parts.append(torch.full((in_dim,), math.log(1.0 / io_scale_lr_factor), dtype=torch.float64))
        parts.append(torch.full((out_dim,), math.log(1.0 / io_scale_lr_factor), dtype=torch.float64))
    return torch.cat(parts).to(dtype)


def default_set_sizes(pde, n):
    """Point-set sizes the reference samplers produce for ``n`` requested points."""
    if pde == "poisson":
        return [n, n]
    if pde == "td_burgers":  # td_burgers_common.py:133-175 with sample_tsteps = 64
        n_scaled = max(n // 63, 1)
        n_sample = 63 * n_scaled
        return [63, 63, n_sample + 2, n_sample]
    if pde == "hyper_elasticity":
        return [n, n, n]
    raise ValueError(pde)


def make_tasks(pde, T, set_sizes, seed=1, dtype=torch.float32):
    """-> (sets: list of [T, n_s, 2], pde_params [T, nprm]) in ABI order."""
    g = torch.Generator().manual_seed(int(seed))
    names = PDE_SETS[pde]
    assert len(set_sizes) == len(names)
    sets = []
    if pde == "poisson":
        nb, nd = set_sizes
        th = torch.rand(T, nb, generator=g, dtype=torch.float64) * 2.0 * math.pi
        sets.append(torch.stack([torch.cos(th), torch.sin(th)], dim=-1))
        # uniform in the unit disc by rejection from [-1,1]^2 (the reference rejects outside r0(theta))
        pts = torch.empty(T, nd, 2, dtype=torch.float64)
        for t in range(T):
            acc = []
            need = nd
            while need > 0:
                c = torch.rand(2 * need + 16, 2, generator=g, dtype=torch.float64) * 2.0 - 1.0
                c = c[(c ** 2).sum(-1) < 1.0][:need]
                acc.append(c)
                need -= c.shape[0]
            pts[t] = torch.cat(acc)[:nd]
        sets.append(pts)
        source = torch.randn(T, 2, 3, generator=g, dtype=torch.float64)
        bc = torch.rand(T, 5, generator=g, dtype=torch.float64) * 2.0 - 1.0
        geo = torch.zeros(T, 2, dtype=torch.float64)
        prm = torch.cat([source.reshape(T, 6), bc, geo], dim=1)
    elif pde == "td_burgers":
        nl, nr, ni, nd = set_sizes
        tl = torch.rand(T, nl, generator=g, dtype=torch.float64)
        sets.append(torch.stack([torch.zeros_like(tl), tl], dim=-1))           # x = xmin = 0
        sets.append(torch.stack([torch.ones(T, nr, dtype=torch.float64), tl[:, :nr] if nr <= nl else
                                 torch.rand(T, nr, generator=g, dtype=torch.float64)], dim=-1))  # x = xmax = 1, same t (quirk 7)
        xi = torch.rand(T, ni, generator=g, dtype=torch.float64)
        sets.append(torch.stack([xi, torch.zeros_like(xi)], dim=-1))           # t = 0
        sets.append(torch.rand(T, nd, 2, generator=g, dtype=torch.float64))
        re = 100.0 * (0.8 + 0.2 * torch.rand(T, 1, generator=g, dtype=torch.float64))
        ic = torch.rand(T, 2, generator=g, dtype=torch.float64) * 4.0 - 2.0
        prm = torch.cat([re, ic], dim=1)
    elif pde == "hyper_elasticity":
        nt_, nb, nd = set_sizes
        xt = torch.rand(T, nt_, generator=g, dtype=torch.float64)
        sets.append(torch.stack([xt, torch.ones_like(xt)], dim=-1))            # y = ymax = 1
        xb = torch.rand(T, nb, generator=g, dtype=torch.float64)
        sets.append(torch.stack([xb, torch.zeros_like(xb)], dim=-1))           # y = ymin = 0
        sets.append(torch.rand(T, nd, 2, generator=g, dtype=torch.float64))
        source = 0.25 + 0.5 * torch.rand(T, 2, generator=g, dtype=torch.float64)
        bc = 0.9 + 0.2 * torch.rand(T, 2, generator=g, dtype=torch.float64)
        prm = torch.cat([source, bc], dim=1)
    else:
        raise ValueError(pde)
    assert prm.shape[1] == PDE_NPRM[pde]
    return [s.to(dtype).contiguous() for s in sets], prm.to(dtype).contiguous()


def out_dim_of(pde):
    return PDE_OUT_DIM[pde]
