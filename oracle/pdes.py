"""Oracle: per-PDE ``loss_fn(field_fn, points, params)`` (TEST INFRASTRUCTURE ONLY).

Mirrors the reference's plugin interface (src/get_pde.py:10-24) and the
autodiff nesting of each loss:
  * Poisson            src/poisson/poisson_common.py:26-39,122-151
  * td-Burgers         src/burgers/td_burgers_common.py:47-91,
                       src/burgers/burgers_formulation/default.py:9-31
  * hyperelasticity    src/elasticity/hyper_elasticity_common.py:23-86
  * driver loss wrapper src/maml_pde.py:76-85 (= leap_pde.py:66-73,
                       nn_pde_maml.py:62-71)
"""
import math

import torch
from torch.func import jacfwd, vmap

from .field import vmap_laplace_operator


# ------------------------------------------------------------------ Poisson
def poisson_boundary_conditions(r, x):
    """poisson_common.py:122-134"""
    theta = torch.atan2(x[1], x[0])
    return (
        r[0]
        + r[1] / 4 * torch.cos(theta)
        + r[2] / 4 * torch.sin(theta)
        + r[3] / 4 * torch.cos(2 * theta)
        + r[4] / 4 * torch.sin(2 * theta)
    ).sum()


def poisson_source(r, x):
    """poisson_common.py:142-146"""
    xx = x.reshape(1, -1) * torch.ones(r.shape[0], x.shape[0], dtype=x.dtype)
    results = r[:, 2] * torch.exp(-((xx[:, 0] - r[:, 0]) ** 2 + (xx[:, 1] - r[:, 1]) ** 2))
    return results.sum()


def poisson_loss_fn(potential_fn, points, params):
    """poisson_common.py:26-39"""
    points_on_boundary, points_in_domain = points
    source_params, bc_params, _ = params
    err_on_boundary = vmap(lambda x: poisson_boundary_conditions(bc_params, x))(points_on_boundary) \
        - potential_fn(points_on_boundary)
    loss_on_boundary = torch.mean(err_on_boundary ** 2)
    err_in_domain = vmap_laplace_operator(
        points_in_domain, potential_fn, lambda x: 1 + 0.1 * potential_fn(x) ** 2
    ) - vmap(lambda x: poisson_source(source_params, x))(points_in_domain)
    loss_in_domain = torch.mean(err_in_domain ** 2)
    return {"boundary_loss": loss_on_boundary}, {"domain_loss": loss_in_domain}


def poisson_residual(potential_fn, points_in_domain, source_params):
    """The per-point domain residual inside poisson_loss_fn (for residual-vector parity)."""
    return vmap_laplace_operator(
        points_in_domain, potential_fn, lambda x: 1 + 0.1 * potential_fn(x) ** 2
    ) - vmap(lambda x: poisson_source(source_params, x))(points_in_domain)


# --------------------------------------------------------------- td-Burgers
def burgers_rhs(field_fn, points_in_domain, params):
    """td_burgers_common.py:54-65 without the final square"""
    source_params, ic_params = params

    def rhs_fn(x):
        jac_fn = jacfwd(field_fn)
        jac_fn_x = lambda x: jac_fn(x)[0]
        hessian_fn = jacfwd(jac_fn_x)
        jac_fn_val = jac_fn(x)
        time_term = jac_fn_val[1]
        hessian = hessian_fn(x)[0]
        nabla_term = (1.0 / source_params[0]) * hessian
        grad_term = jac_fn_val[0] * field_fn(x)
        return time_term - (nabla_term - grad_term)

    return vmap(rhs_fn)(points_in_domain)


def burgers_loss_domain_fn(field_fn, points_in_domain, params):
    """td_burgers_common.py:47-67"""
    return burgers_rhs(field_fn, points_in_domain, params) ** 2


def burgers_loss_initial_fn(field_fn, points_on_initial, params):
    """burgers_formulation/default.py:9-23"""
    source_params, ic_params = params
    x = points_on_initial[:, 0]
    sinusoidal_magnitude_x = (
        torch.sin(math.pi * x)
        + ic_params[0] * torch.sin(2.0 * math.pi * x)
        + ic_params[1] * torch.sin(4.0 * math.pi * x)
    )
    return (field_fn(points_on_initial) - sinusoidal_magnitude_x) ** 2


def burgers_loss_fn(field_fn, points, params):
    """td_burgers_common.py:70-91 (left/right reuse the IC formula, default.py:26-31)"""
    points_on_left, points_on_right, points_initial, points_in_domain = points
    return (
        {
            "loss_initial": torch.mean(burgers_loss_initial_fn(field_fn, points_initial, params)),
            "loss_left": torch.mean(burgers_loss_initial_fn(field_fn, points_on_left, params)),
            "loss_right": torch.mean(burgers_loss_initial_fn(field_fn, points_on_right, params)),
        },
        {"loss_domain": torch.mean(burgers_loss_domain_fn(field_fn, points_in_domain, params))},
    )


# ---------------------------------------------------------- hyperelasticity
def deformation_gradient(x, field_fn):
    """hyper_elasticity_common.py:23-26"""
    jac = jacfwd(lambda x: field_fn(x).squeeze())(x)
    return torch.eye(2, dtype=x.dtype) + jac


def right_cauchygreen(x, field_fn):
    """hyper_elasticity_common.py:29-31"""
    F = deformation_gradient(x, field_fn)
    return torch.matmul(F, F.transpose(0, 1))


def hyper_energy_density(field_fn, points_in_domain, params):
    """hyper_elasticity_common.py:34-56"""
    source_params, bc_params, per_hole_params, n_holes = params

    def integrand(x):
        young_mod = bc_params[0]
        poisson_ratio = 0.49
        d = 2
        shear_mod = young_mod / (2 * (1 + poisson_ratio))
        bulk_mod = young_mod / (3 * (1 - 2 * poisson_ratio))
        F = deformation_gradient(x, field_fn)
        J = F[0, 0] * F[1, 1] - F[0, 1] * F[1, 0]  # np.linalg.det of a 2x2
        Jinv = J ** (-2 / d)
        Ic = torch.trace(right_cauchygreen(x, field_fn))
        return (shear_mod / 2) * (Jinv * Ic - d) + (bulk_mod / 2) * (J - 1) ** 2

    return vmap(integrand)(points_in_domain)


def hyper_loss_fn(field_fn, points, params):
    """hyper_elasticity_common.py:59-86"""
    (points_on_top, points_on_bottom, points_on_left, points_on_right,
     points_on_holes, points_in_domain) = points
    top_target = torch.tensor([0.0, -0.12], dtype=points_on_top.dtype)
    loss_top = (field_fn(points_on_top) - top_target) ** 2
    loss_bottom = field_fn(points_on_bottom) ** 2
    return (
        {
            "loss_bottom": 1000.0 * torch.mean(loss_bottom),
            "loss_top": 1000.0 * torch.mean(loss_top),
        },
        {"loss_domain": torch.mean(hyper_energy_density(field_fn, points_in_domain, params))},
    )


# ------------------------------------------------------------------ registry
PDE_LOSS = {
    "poisson": poisson_loss_fn,
    "td_burgers": burgers_loss_fn,
    "hyper_elasticity": hyper_loss_fn,
}
PDE_OUT_DIM = {"poisson": 1, "td_burgers": 1, "hyper_elasticity": 2}
# order of the aux dict {**boundary_losses, **domain_losses} (maml_pde.py:85)
PDE_TERMS = {
    "poisson": ["boundary_loss", "domain_loss"],
    "td_burgers": ["loss_initial", "loss_left", "loss_right", "loss_domain"],
    "hyper_elasticity": ["loss_bottom", "loss_top", "loss_domain"],
}


def driver_loss(pde_name, field_fn, points, params, bc_weight):
    """maml_pde.py:76-85: loss = bc_weight * sum(boundary) + sum(domain); aux = merged dict."""
    boundary_losses, domain_losses = PDE_LOSS[pde_name](field_fn, points, params)
    loss = bc_weight * torch.sum(torch.stack([bl for bl in boundary_losses.values()])) + torch.sum(
        torch.stack([dl for dl in domain_losses.values()])
    )
    return loss, {**boundary_losses, **domain_losses}
