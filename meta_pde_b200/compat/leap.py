"""LEAP on the plugin surface: same entry points as /root/reference/src/nets/leap.py.

    LeapDef                                                     leap.py:32-47
    leap_inner_step / single_task_rollout / single_task_grad_and_losses / multi_task_grad_and_losses   :49-187
    get_meta_grad_increment / compute_global_norm               :190-228

Fused route (``compat.tasks.PdeTaskLoss`` losses + Adam or Momentum(beta=0) inner optimizer): one call of
mpde_leap_outer for all tasks.  Generic route: the reference's algorithm over the kernel-backed loss, one task at a time.
"""
from collections import namedtuple

import torch

from .. import ops
from . import random as jr
from .field import Model
from .flags import FLAGS
from .optim import Adam, Momentum
from .tasks import PdeTaskLoss

LeapDef = namedtuple("LeapDef", ["make_inner_opt", "make_task_loss_fn", "inner_steps", "n_batch_tasks", "norm",
                                 "loss_in_distance", "stabilize"])


def compute_global_norm(leap_def, new_model, old_model, d_loss):
    sum_sq = torch.sum((new_model.theta - old_model.theta) ** 2)
    if leap_def.loss_in_distance:
        sum_sq = sum_sq + d_loss ** 2
    return torch.sqrt(sum_sq)


def get_meta_grad_increment(leap_def, new_model, model, new_loss, loss, grad):
    d_loss = new_loss - loss
    if leap_def.stabilize:
        d_loss = -torch.abs(d_loss)
    norm = compute_global_norm(leap_def, new_model, model, d_loss) if leap_def.norm else 1.0
    inc = model.theta - new_model.theta
    if leap_def.loss_in_distance:
        inc = inc - d_loss * grad.theta
    return Model(model.module, inc / norm)


def leap_inner_step(leap_def, key, opt, loss_fn, meta_grad_accum):
    k1, k2 = jr.split(key, 2)
    theta = opt.target.theta.detach().requires_grad_(True)
    loss, aux = loss_fn(k1, Model(opt.target.module, theta))
    (g,) = torch.autograd.grad(loss, theta)
    loss = loss.detach()
    grad_norm = torch.sqrt(torch.sum(g ** 2))
    if bool(grad_norm > FLAGS.inner_grad_clip):
        g = FLAGS.inner_grad_clip * g / grad_norm
    grad = Model(opt.target.module, g)
    new_opt = opt.apply_gradient(grad)
    new_loss, new_aux = loss_fn(k2, new_opt.target)
    inc = get_meta_grad_increment(leap_def, new_opt.target, opt.target, new_loss.detach(), loss, grad)
    return new_opt, meta_grad_accum.tree_map(lambda a, b: a + b, inc), (new_loss.detach(), new_aux)


def _rollout_keys(key, K):
    loss0_key, inner_key = jr.split(key, 2)
    return loss0_key, jr.split(inner_key, K)


def single_task_rollout(leap_def, key, initial_model, loss_fn):
    """leap.py:105-147 -> (final_model, meta_grad_accum, (losses [K+1], aux))."""
    loss0_key, inner_keys = _rollout_keys(key, leap_def.inner_steps)
    loss0, aux0 = loss_fn(loss0_key, initial_model)
    opt = leap_def.make_inner_opt(Model(initial_model.module, initial_model.theta.detach()))
    accum = Model(initial_model.module, torch.zeros_like(initial_model.theta))
    losses = [loss0.detach()]
    for k in range(leap_def.inner_steps):
        opt, accum, (new_loss, _) = leap_inner_step(leap_def, inner_keys[k], opt, loss_fn, accum)
        losses.append(new_loss)
    return opt.target, accum, torch.stack(losses)


def single_task_grad_and_losses(leap_def, key, initial_model):
    loss_fn_key, rollout_key = jr.split(key, 2)
    loss_fn = leap_def.make_task_loss_fn(loss_fn_key)
    _, meta_grad, losses = single_task_rollout(leap_def, rollout_key, initial_model, loss_fn)
    return meta_grad, losses


def multi_task_grad_and_losses(leap_def, key, initial_model, task_slice=None, grad_scale=None):
    """leap.py:170-187 -> (grads Model, losses [T, K+1]).  task_slice / grad_scale: as in compat.maml."""
    T_all, K = leap_def.n_batch_tasks, leap_def.inner_steps
    keys = jr.split(key, T_all)
    lo, hi = task_slice if task_slice is not None else (0, T_all)
    scale = (1.0 / T_all) if grad_scale is None else grad_scale
    od = getattr(leap_def.make_inner_opt, "__self__", None)
    tasks = []
    for t in range(lo, hi):
        loss_fn_key, rollout_key = jr.split(keys[t], 2)
        tasks.append((leap_def.make_task_loss_fn(loss_fn_key), rollout_key))
    if isinstance(od, (Adam, Momentum)) and all(isinstance(f, PdeTaskLoss) for f, _ in tasks):
        return _fused(leap_def, od, tasks, initial_model, K, scale)
    outs = [single_task_rollout(leap_def, rk, initial_model, f) for f, rk in tasks]
    return (Model(initial_model.module, scale * torch.stack([o[1].theta for o in outs]).sum(0)),
            torch.stack([o[2] for o in outs]))


def _fused(leap_def, od, tasks, initial_model, K, scale):
    from .maml import _stack_points

    dev = initial_model.theta.device
    K_ = ops.kernels(dev)
    pde = tasks[0][0].pde
    prms = [pde.abi_params(fn.params) for fn, _ in tasks]
    sets = None
    if FLAGS.gpu_samplers:                                   # all (2K + 1) x T draws in one launch, same keys (compat/gpu_samplers.py)
        import numpy as np
        from . import gpu_samplers
        T = len(tasks)
        keys = np.zeros((2 * K + 1, T, 2), np.uint32)
        for t, (fn, rollout_key) in enumerate(tasks):
            loss0_key, inner_keys = _rollout_keys(rollout_key, K)
            keys[0, t] = loss0_key
            for k in range(K):
                keys[1 + k, t], keys[1 + K + k, t] = jr.split(inner_keys[k], 2)
        got = gpu_samplers.sample_points_batched(pde, keys.reshape(-1, 2), tasks[0][0].n_points, [fn.params for fn, _ in tasks] * (2 * K + 1), K_)
        if got is not None:
            pick = lambda i: [x.reshape(2 * K + 1, T, *x.shape[1:])[i].contiguous() for x in got]
            sets = (pick(0), [pick(1 + k) for k in range(K)], [pick(1 + K + k) for k in range(K)])
    if sets is None:
        loss0, step, new = [], [[] for _ in range(K)], [[] for _ in range(K)]
        for fn, rollout_key in tasks:
            loss0_key, inner_keys = _rollout_keys(rollout_key, K)
            loss0.append(fn.sample(loss0_key))
            for k in range(K):
                k1, k2 = jr.split(inner_keys[k], 2)
                step[k].append(fn.sample(k1)); new[k].append(fn.sample(k2))
        sets = (_stack_points(pde, loss0), [_stack_points(pde, p) for p in step], [_stack_points(pde, p) for p in new])
    cfg = dict(inner_steps=K, inner_lr=od.learning_rate, inner_grad_clip=FLAGS.inner_grad_clip, norm=leap_def.norm,
               loss_in_distance=leap_def.loss_in_distance, stabilize=leap_def.stabilize,
               inner_opt="adam" if isinstance(od, Adam) else "sgd", beta1=getattr(od, "beta1", 0.9),
               beta2=getattr(od, "beta2", 0.99), eps=getattr(od, "eps", 1e-8))
    out = K_.leap_outer(initial_model.net_desc(K_), pde.name, tasks[0][0].bc_weight, cfg, initial_model.theta.detach().contiguous(),
                        sets[0], sets[1], sets[2],
                        torch.stack(prms).to(dev).contiguous(), len(tasks), grad_scale=scale)
    return Model(initial_model.module, out["meta_grad"]), out["losses"]
