"""Deployment / validation rollout on the plugin surface.

Mirrors the evaluation closures the reference's drivers define inside ``main`` and the FEniCS-free half of its
validation:

    get_final_model, make_coef_func, partial_make_coef_func, validation_losses   src/maml_pde.py:163-201
    the LEAP twins                                                               src/leap_pde.py:140-173
    vmap_validation_error, vmap_validation_rollout                               src/util/trainer_util.py:501-590

``get_final_model`` adapts one task: fresh inner points from ``k1`` (held FIXED over the inner steps, as the reference's
closure ignores the per-step key), ``inner_steps`` SGD steps with the learned per-parameter learning rates through
``mpde_maml_adapt``; ``make_coef_func`` evaluates the adapted field with ``mpde_field_eval``.  The reference vmaps
``make_coef_func`` over the ``n_eval`` validation tasks; here the closures carry a ``batched`` twin that adapts ALL tasks
in one ``mpde_maml_adapt`` call (grid.y = task) and evaluates them in one ``mpde_field_eval`` call, and
``vmap_validation_error`` uses it when present.  Ground-truth values are an INPUT here: producing them is FEniCS's job
(out of scope); the error arithmetic, including the hyper-elasticity mirror quirk
(trainer_util.py:524-547 keeps ``mse_left`` in both branches) and the per-time-step Burgers error, is restated.
"""
from types import SimpleNamespace

import numpy as np
import torch

from .. import ops
from . import leap as cleap
from . import maml as cmaml
from . import random as jr
from .field import Model
from .flags import FLAGS
from .tasks import PdeTaskLoss, driver_loss


class FixedPointsLoss(PdeTaskLoss):
    """``lambda key, field_fn: loss_fn(field_fn, inner_points, params)`` of maml_pde.py:169 as a recognisable object."""

    def __init__(self, pde, params, points, bc_weight=None):
        super().__init__(pde, params, 0, bc_weight)
        self.points = points

    def sample(self, key):
        return self.points

    def __call__(self, key, field_fn):
        return driver_loss(self.pde, field_fn, self.points, self.params, self.bc_weight)


def _unstack_params(params, i):
    """vmapped ``sample_params`` output (tuple of arrays with a leading task axis) or a list of per-task tuples -> task i."""
    if isinstance(params, (list,)) or (isinstance(params, tuple) and len(params) and isinstance(params[0], tuple)):
        return params[i]
    return tuple(p[i] if (torch.is_tensor(p) or isinstance(p, np.ndarray)) and getattr(p, "ndim", 0) > 0 else p for p in params)


def _n_tasks(params):
    if isinstance(params, list) or (isinstance(params, tuple) and len(params) and isinstance(params[0], tuple)):
        return len(params)
    return int(params[0].shape[0])


def make_maml_eval_fns(pde, maml_def):
    """-> namespace(get_final_model, make_coef_func, partial_make_coef_func, validation_losses)  (maml_pde.py:163-201)"""

    def get_final_model(key, model_and_lrs, params, inner_steps, maml_def=maml_def):
        model, inner_lrs = model_and_lrs
        k1, k2 = jr.split(key, 2)
        inner_points = pde.sample_points(k1, FLAGS.inner_points, params)
        inner_loss_fn = FixedPointsLoss(pde, params, inner_points)
        if inner_steps == 0:
            return model
        lrs = Model(inner_lrs.module, inner_lrs.theta[:inner_steps]) if inner_lrs is not None else None
        return cmaml.single_task_rollout(maml_def._replace(inner_steps=inner_steps), k2, model, inner_loss_fn, lrs)[0]

    def make_coef_func(key, model_and_lrs, params, coords, inner_steps, maml_def=maml_def):
        final_model = get_final_model(key, model_and_lrs, params, inner_steps, maml_def)
        return torch.squeeze(final_model(coords))

    def get_final_models(keys, model_and_lrs, params, inner_steps, maml_def=maml_def):
        """All tasks at once: one mpde_maml_adapt call.  -> theta_final [T, P]"""
        model, inner_lrs = model_and_lrs
        dev = model.theta.device
        T = len(keys)
        if inner_steps == 0:
            return model.theta.detach().unsqueeze(0).repeat(T, 1).contiguous()
        K_ = ops.kernels(dev)
        pts, prms = [], []
        for i in range(T):
            k1, _ = jr.split(keys[i], 2)
            p_i = _unstack_params(params, i)
            pts.append(pde.sample_points(k1, FLAGS.inner_points, p_i))
            prms.append(pde.abi_params(p_i))
        sets = cmaml._stack_points(pde, pts)
        prm = torch.stack(prms).to(dev).contiguous()
        lrs = inner_lrs.theta[:inner_steps].contiguous() if inner_lrs is not None else None
        thf, _ = K_.maml_adapt(model.net_desc(K_), pde.name, float(FLAGS.bc_weight), cmaml._cfg(maml_def, inner_steps),
                               model.theta.detach().contiguous(), lrs, [sets] * inner_steps, sets, prm, T)
        return thf

    def make_coef_funcs(keys, model_and_lrs, params, coords, inner_steps=None, maml_def=maml_def):
        """vmap(make_coef_func, (0, None, 0, 0)) of trainer_util.py:508 as two launches.  coords [T, n, 2] -> [T, n, out]"""
        model = model_and_lrs[0]
        steps = maml_def.inner_steps if inner_steps is None else inner_steps
        thf = get_final_models(keys, model_and_lrs, params, steps, maml_def)
        K_ = ops.kernels(model.theta.device)
        c = torch.as_tensor(coords, dtype=torch.float32, device=model.theta.device).contiguous()
        return K_.field_eval(model.net_desc(K_), thf, c, len(keys))

    def partial_make_coef_func(key, model_and_lrs, params, coords):
        return make_coef_func(key, model_and_lrs, params, coords, maml_def.inner_steps, maml_def)
    partial_make_coef_func.batched = lambda keys, model_and_lrs, params, coords: make_coef_funcs(keys, model_and_lrs, params, coords)

    def validation_losses(model_and_lrs, maml_def=maml_def):
        model, inner_lrs = model_and_lrs
        _, losses, meta_losses = cmaml.multi_task_grad_and_losses(maml_def, jr.PRNGKey(0), model, inner_lrs)
        return losses, meta_losses

    return SimpleNamespace(get_final_model=get_final_model, make_coef_func=make_coef_func, get_final_models=get_final_models,
                           make_coef_funcs=make_coef_funcs, partial_make_coef_func=partial_make_coef_func,
                           validation_losses=validation_losses)


def make_leap_eval_fns(pde, leap_def):
    """-> namespace(get_final_model, make_coef_func, partial_make_coef_func, validation_losses)  (leap_pde.py:140-173)"""

    def get_final_model(key, model, params, inner_steps, leap_def=leap_def):
        k1, k2 = jr.split(key, 2)
        inner_points = pde.sample_points(k1, FLAGS.inner_points, params)
        inner_loss_fn = FixedPointsLoss(pde, params, inner_points)
        if inner_steps == 0:
            return model
        return cleap.single_task_rollout(leap_def._replace(inner_steps=inner_steps), k2, model, inner_loss_fn)[0]

    def make_coef_func(key, model, params, coords, inner_steps, leap_def=leap_def):
        return torch.squeeze(get_final_model(key, model, params, inner_steps, leap_def)(coords))

    def partial_make_coef_func(key, model, params, coords):
        return make_coef_func(key, model, params, coords, leap_def.inner_steps, leap_def)

    def validation_losses(model, leap_def=leap_def):
        _, losses = cleap.multi_task_grad_and_losses(leap_def, jr.PRNGKey(0), model)
        return losses

    return SimpleNamespace(get_final_model=get_final_model, make_coef_func=make_coef_func,
                           partial_make_coef_func=partial_make_coef_func, validation_losses=validation_losses)


# --------------------------------------------------------------------------------------------- trainer_util.py:501-590
def vmap_validation_rollout(model, ground_truth_params, points, ground_truth_vals, make_coef_func):
    """trainer_util.py:578-590: the adapted fields of the n_eval validation tasks at `points` [n_eval, n, 2]."""
    keys = jr.split(jr.PRNGKey(0), FLAGS.n_eval)
    pts = torch.as_tensor(points, dtype=torch.float32)
    if hasattr(make_coef_func, "batched"):
        return make_coef_func.batched(keys, model, ground_truth_params, pts)
    dev = (model[0] if isinstance(model, tuple) else model).theta.device
    out = [make_coef_func(keys[i], model, _unstack_params(ground_truth_params, i), pts[i].to(dev)) for i in range(FLAGS.n_eval)]
    return torch.stack(out)


def vmap_validation_error(model, ground_truth_params, points, ground_truth_vals, make_coef_func):
    """trainer_util.py:501-575 -> (mse, mean normalizer, mean rel_sq_err, per-dim rel_sq_err, std over tasks, per-t err)."""
    assert _n_tasks(ground_truth_params) == FLAGS.n_eval
    coefs = vmap_validation_rollout(model, ground_truth_params, points, ground_truth_vals, make_coef_func)
    coefs = coefs.reshape(coefs.shape[0], coefs.shape[1], -1).double()
    gt = torch.as_tensor(ground_truth_vals, dtype=torch.float64, device=coefs.device).reshape(coefs.shape)
    if FLAGS.pde == "hyper_elasticity":
        err_left = coefs - gt
        mse_left = (err_left ** 2).mean(dim=(1, 2))
        pts_r = torch.as_tensor(points, dtype=torch.float32).clone()
        pts_r[:, :, 0] = 1.0 - pts_r[:, :, 0]
        coefs_r = vmap_validation_rollout(model, ground_truth_params, pts_r, ground_truth_vals, make_coef_func)
        coefs_r = coefs_r.reshape(coefs.shape).double().clone()
        coefs_r[:, :, 0] *= -1.0
        err_right = coefs_r - gt
        mse_right = (err_right ** 2).mean(dim=(1, 2))
        take_right = (mse_left > mse_right).reshape(-1, 1, 1)
        err = torch.where(take_right, err_right, err_left)
        mse = mse_left.sum()                       # the reference returns mse_left in BOTH branches (:533-540)
    else:
        err = coefs - gt
        mse = (err ** 2).mean()
    normalizer = (gt ** 2).mean(dim=1, keepdim=True)
    rel_sq_err = err ** 2 / normalizer.mean(dim=2, keepdim=True)
    t_rel = None
    if FLAGS.pde == "td_burgers":
        assert points.shape[-1] == 2
        tile_idx = points.shape[1] // FLAGS.num_tsteps
        t_idx = torch.arange(0, tile_idx, device=coefs.device) * FLAGS.num_tsteps
        t_rel = []
        for i in range(FLAGS.num_tsteps):
            t_err = err[:, t_idx + i, :]
            t_norm = (gt[:, t_idx + i, :] ** 2).mean(dim=1, keepdim=True)
            t_rel.append((t_err ** 2 / t_norm.mean(dim=2, keepdim=True)).mean())
        t_rel = torch.stack(t_rel)
    return (mse, normalizer.mean(dim=(0, 1)), rel_sq_err.mean(), rel_sq_err.mean(dim=(0, 1)),
            rel_sq_err.mean(dim=(1, 2)).std(unbiased=False), t_rel)
