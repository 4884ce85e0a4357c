"""Task-loss closures of the drivers (src/maml_pde.py:76-98, src/leap_pde.py:66-85, src/nn_pde_maml.py:62-78) as
objects, so that the meta-learning entry points can recognise them and batch all tasks into the fused kernels."""
import torch

from .flags import FLAGS


def driver_loss(pde, field_fn, points, params, bc_weight=None):
    """loss = bc_weight * sum(boundary terms) + sum(domain terms); aux = merged dict (src/maml_pde.py:76-85)."""
    bw = FLAGS.bc_weight if bc_weight is None else bc_weight
    boundary_losses, domain_losses = pde.loss_fn(field_fn, points, params)
    loss = bw * torch.stack(list(boundary_losses.values())).sum() + torch.stack(list(domain_losses.values())).sum()
    return loss, {**boundary_losses, **domain_losses}


class PdeTaskLoss:
    """``loss(key, field_fn) -> (loss, aux)`` for one task: fresh points from ``key`` at every call."""

    def __init__(self, pde, params, n_points, bc_weight=None):
        self.pde, self.params, self.n_points = pde, params, int(n_points)
        self.bc_weight = float(FLAGS.bc_weight if bc_weight is None else bc_weight)

    def sample(self, key):
        return self.pde.sample_points(key, self.n_points, self.params)

    def __call__(self, key, field_fn):
        return driver_loss(self.pde, field_fn, self.sample(key), self.params, self.bc_weight)


def make_task_loss_fns_factory(pde, inner_points=None, outer_points=None):
    """-> make_task_loss_fns(key) as in src/maml_pde.py:87-98."""
    def make_task_loss_fns(key):
        params = pde.sample_params(key)
        return (PdeTaskLoss(pde, params, FLAGS.inner_points if inner_points is None else inner_points),
                PdeTaskLoss(pde, params, FLAGS.outer_points if outer_points is None else outer_points))
    return make_task_loss_fns


def make_task_loss_fn_factory(pde, inner_points=None):
    """-> make_task_loss_fn(key) as in src/leap_pde.py:75-85."""
    def make_task_loss_fn(key):
        return PdeTaskLoss(pde, pde.sample_params(key), FLAGS.inner_points if inner_points is None else inner_points)
    return make_task_loss_fn
