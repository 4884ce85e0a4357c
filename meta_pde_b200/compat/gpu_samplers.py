"""Collocation-point samplers on the GPU: the reference's ``sample_points`` for MANY (task, draw)
samples in one launch, with jax 0.2.20's threefry PRNG restated in ``csrc/sampler_kernels.cuh`` (bit-exact with
``compat/random.py``).  Same key in, same points out as the host samplers of ``compat/pdes.py`` (which are pinned to the
reference's sampler source, ``tests/test_sampler_parity.py``):

    poisson      src/poisson/poisson_common.py:69-119      boundary + domain (rejection by Gumbel top-k, the arctan2 quirk)
    td_burgers   src/burgers/td_burgers_common.py:117-186  left / right (shared key: quirk 7) / initial / domain

Uniform draws and key splits are bit-exact.  Transcendentals (cos / sin / atan2 / log) are the CUDA single-precision ones:
boundary points agree with the host samplers to ~1e-7, and the domain selection picks the same candidates except where two
Gumbel keys (or a hole-boundary test) differ by an ulp -- the same latitude a real JAX run has against numpy.
The n-smallest selection of ``jax.random.choice(replace=False, p)`` is a stable device sort (torch.sort: data preparation,
not the hot path).  Hyper-elasticity (a data-dependent ``while_loop`` over pore layouts) stays on the host samplers.
"""
import numpy as np
import torch

from .. import ops
from .flags import FLAGS


def _kernels(device=None):
    return ops.kernels(device if device is not None else torch.device("cuda", torch.cuda.current_device()))


def poisson_sample_points_batched(keys, n, geo_params, K=None):
    """keys [S, 2] uint32, geo_params [S, 2] -> (points_on_boundary [S, n, 2], points_in_domain [S, n, 2]) on K.device."""
    K = K or _kernels()
    boundary, cand_xy, cand_g = K.sample_poisson_candidates(keys, n, geo_params, FLAGS.xmin, FLAGS.xmax, FLAGS.ymin, FLAGS.ymax)
    idx = torch.sort(cand_g, dim=1, stable=True).indices[:, :n]          # argsort(g)[:n]  (compat/random.py: kind="stable")
    domain = torch.gather(cand_xy, 1, idx.unsqueeze(-1).expand(-1, -1, 2)).contiguous()
    return boundary, domain


def _col(x, val):
    return torch.stack([torch.full_like(x, val), x], dim=-1)


def burgers_sample_points_batched(keys, n, K=None):
    """keys [S, 2] uint32 -> (left [S, nt, 2], right [S, nt, 2], initial [S, n_s + 2, 2], domain [S, n_s, 2])."""
    if not FLAGS.sample_time_random:
        raise NotImplementedError("gpu samplers: td_burgers with --nosample_time_random uses the host sampler")
    K = K or _kernels()
    nt = FLAGS.sample_tsteps - 1
    k = K.random_split(keys, 5)                                          # k1 .. k5 (td_burgers_common.py:119)
    k2, k3, k5 = k[:, 1].contiguous(), k[:, 2].contiguous(), k[:, 4].contiguous()
    kt = K.random_split(k2, 2)[:, 1].contiguous()                        # left and right share k2 (:120-121, :135, :147)
    t_side = K.random_uniform(kt, nt, FLAGS.tmin, FLAGS.tmax)            # sample_time(k2', 1)
    left, right = _col(t_side, float(FLAGS.xmin)), _col(t_side, float(FLAGS.xmax))

    def in_domain(kx):                                                   # :158-169
        kk = K.random_split(kx, 2)
        n_scaled = n // nt
        xs = K.random_uniform(kk[:, 0].contiguous(), n_scaled * nt, FLAGS.xmin, FLAGS.xmax)
        ts = K.random_uniform(kk[:, 1].contiguous(), nt * n_scaled, FLAGS.tmin, FLAGS.tmax)
        return torch.stack([xs, ts], dim=-1)

    dom_i = in_domain(k3)                                                # sample_points_initial (:172-176)
    S = dom_i.shape[0]
    ends = torch.tensor([float(FLAGS.xmin), float(FLAGS.xmax)], device=dom_i.device).expand(S, 2)
    xi = torch.cat([dom_i[:, :, 0], ends], dim=1)
    initial = torch.stack([xi, torch.zeros_like(xi)], dim=-1)
    return left.contiguous(), right.contiguous(), initial.contiguous(), in_domain(k5).contiguous()


def sample_points_batched(pde, keys, n, params_list, K=None):
    """ABI point sets [S, n_s, 2] of ``pde.sample_points(keys[s], n, params_list[s])`` for all s, or None when this PDE has
    no GPU sampler (the caller falls back to the host sampler)."""
    keys = np.ascontiguousarray(np.asarray(keys, dtype=np.uint32).reshape(-1, 2))
    if pde.name == "poisson":
        geo = torch.stack([torch.as_tensor(p[2], dtype=torch.float32).reshape(2).cpu() for p in params_list])
        return list(poisson_sample_points_batched(keys, n, geo, K))
    if pde.name == "td_burgers" and FLAGS.sample_time_random:
        return list(burgers_sample_points_batched(keys, n, K))
    return None
