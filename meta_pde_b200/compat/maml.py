"""MAML on the plugin surface: same entry points and return structures as /root/reference/src/nets/maml.py.

    MamlDef                                   maml.py:31-47
    maml_inner_step(key, opt, inner_loss_fn, inner_lr, softplus_lrs)            :50-97
    single_task_rollout(maml_def, rollout_key, initial_model, inner_loss_fn, inner_lrs, inner_steps, outer_loss_fn)  :100-166
    single_task_grad_and_losses(maml_def, key, initial_model, inner_lrs)        :169-207
    multi_task_grad_and_losses(maml_def, key, initial_model, inner_lrs)         :210-233

Two execution routes, both on the GPU kernels (there is no CPU route):
  * fused: when the task losses are ``compat.tasks.PdeTaskLoss`` objects and the inner optimizer is Momentum(beta=0),
    all tasks' points are sampled up front (same key derivations as the reference's scan) and ONE call of
    mpde_maml_outer / mpde_maml_adapt does the forward sweep and the reverse sweep of Hessian-vector products;
  * generic: any loss closure built on ``pde.loss_fn`` with a SIREN Model: the reference's algorithm written with
    torch autograd, whose nodes are the fused loss / adjoint / HVP kernels (one task at a time).
"""
from collections import namedtuple

import torch
import torch.nn.functional as F

from .. import ops
from .._abi import PDE_TERMS
from . import random as jr
from .field import Model
from .flags import FLAGS
from .optim import Momentum
from .tasks import PdeTaskLoss

MamlDef = namedtuple("MamlDef", ["make_inner_opt", "make_task_loss_fns", "inner_steps", "n_batch_tasks", "softplus_lrs",
                                 "outer_loss_decay"])


def _inner_opt_def(maml_def):
    return getattr(maml_def.make_inner_opt, "__self__", None)


def _fusable(maml_def, fns):
    od = _inner_opt_def(maml_def)
    return isinstance(od, Momentum) and all(isinstance(f, PdeTaskLoss) for f in fns if f is not None)


def _stack_points(pde, tuples):
    """[task] reference point tuples -> ABI sets, each [T, n, 2]"""
    per_task = [pde.abi_sets(p) for p in tuples]
    return [torch.stack([pt[s] for pt in per_task]).contiguous() for s in range(len(per_task[0]))]


def _rollout_keys(rollout_key, K):
    """The key chain of the scan body (maml.py:132-140): per step (k1, k2), then the final key."""
    key, k1s, k2s = rollout_key, [], []
    for _ in range(K):
        k1, k2, key = jr.split(key, 3)
        k1s.append(k1); k2s.append(k2)
    return k1s, k2s, key


def _cfg(maml_def, K):
    od = _inner_opt_def(maml_def)
    return dict(inner_steps=K, inner_lr=od.learning_rate, inner_grad_clip=FLAGS.inner_grad_clip,
                softplus_lrs=bool(maml_def.softplus_lrs), outer_loss_decay=float(maml_def.outer_loss_decay or 0.0))


# ---------------------------------------------------------------------------------------------- generic route
def maml_inner_step(key, opt, inner_loss_fn, inner_lr, softplus_lrs=False):
    """maml.py:50-97 with torch autograd over the kernel-backed loss."""
    model = opt.target
    diff = model.theta.requires_grad          # inside an outer differentiation (second-order route)?
    theta = model.theta if diff else model.theta.detach().requires_grad_(True)
    loss, _ = inner_loss_fn(key, Model(model.module, theta))
    (grad,) = torch.autograd.grad(loss, theta, create_graph=diff)
    if not diff:
        theta, loss = theta.detach(), loss.detach()
    sp = (lambda x: F.softplus(x)) if softplus_lrs else (lambda x: x)
    lr = inner_lr.theta if isinstance(inner_lr, Model) else torch.as_tensor(float(inner_lr), device=theta.device)
    grad = grad * sp(lr)
    grad_norm = torch.sqrt(torch.sum(grad ** 2))
    if bool(grad_norm > FLAGS.inner_grad_clip):
        grad = FLAGS.inner_grad_clip * grad / grad_norm
    new_opt = opt.optimizer_def.create(Model(model.module, theta)).apply_gradient(Model(model.module, grad))
    return new_opt, loss


def single_task_rollout(maml_def, rollout_key, initial_model, inner_loss_fn, inner_lrs=None, inner_steps=-1,
                        outer_loss_fn=None):
    """maml.py:100-166 -> (final_model, (meta_loss_sum, losses [K+1]))."""
    if inner_lrs is not None:
        assert inner_steps == -1
    K = maml_def.inner_steps if inner_steps < 0 else inner_steps
    if outer_loss_fn is None and _fusable(maml_def, [inner_loss_fn]) and not initial_model.theta.requires_grad:
        return _fused_adapt(maml_def, rollout_key, initial_model, inner_loss_fn, inner_lrs, K)
    k1s, k2s, final_key = _rollout_keys(rollout_key, K)
    opt = maml_def.make_inner_opt(initial_model)
    meta_loss, losses = torch.zeros((), device=initial_model.theta.device), []
    for k in range(K):
        lr = Model(inner_lrs.module, inner_lrs.theta[k]) if isinstance(inner_lrs, Model) else 1.0
        opt, loss = maml_inner_step(k1s[k], opt, inner_loss_fn, lr, maml_def.softplus_lrs)
        losses.append(loss)
        if outer_loss_fn is not None:
            meta_loss = outer_loss_fn(k2s[k], opt.target)[0] + meta_loss * maml_def.outer_loss_decay
    loss_final, _ = inner_loss_fn(final_key, opt.target)
    return opt.target, (meta_loss, torch.stack(losses + [loss_final]))


def _fused_adapt(maml_def, rollout_key, initial_model, inner_loss_fn, inner_lrs, K):
    K_ = ops.kernels(initial_model.theta.device)
    pde = inner_loss_fn.pde
    k1s, _, final_key = _rollout_keys(rollout_key, K)
    inner = [_stack_points(pde, [inner_loss_fn.sample(k)]) for k in k1s]
    final = _stack_points(pde, [inner_loss_fn.sample(final_key)])
    prm = pde.abi_params(inner_loss_fn.params).reshape(1, -1).to(initial_model.theta.device).contiguous()
    lrs = inner_lrs.theta[:K].contiguous() if isinstance(inner_lrs, Model) else None
    thf, losses = K_.maml_adapt(initial_model.net_desc(K_), pde.name, inner_loss_fn.bc_weight, _cfg(maml_def, K),
                                initial_model.theta.contiguous(), lrs, inner, final, prm, 1)
    return Model(initial_model.module, thf[0]), (torch.zeros((), device=thf.device), losses[0])


def single_task_grad_and_losses(maml_def, key, initial_model, inner_lrs=None):
    """maml.py:169-207 (generic route) -> (meta_grad, losses, (meta_loss, outer_aux))."""
    task_key, rollout_key, outer_loss_key = jr.split(key, 3)
    inner_loss_fn, outer_loss_fn = maml_def.make_task_loss_fns(task_key)
    theta = initial_model.theta.detach().clone().requires_grad_(True)
    lrs = inner_lrs.theta.detach().clone().requires_grad_(True) if inner_lrs is not None else None
    final_model, (outer_loss, losses) = single_task_rollout(
        maml_def, rollout_key, Model(initial_model.module, theta), inner_loss_fn,
        Model(inner_lrs.module, lrs) if lrs is not None else None, outer_loss_fn=outer_loss_fn)
    _, outer_aux = outer_loss_fn(outer_loss_key, Model(final_model.module, final_model.theta.detach()))
    wrt = (theta, lrs) if lrs is not None else (theta,)
    g = torch.autograd.grad(outer_loss, wrt, allow_unused=True)
    g = [gi if gi is not None else torch.zeros_like(w) for gi, w in zip(g, wrt)]
    meta_grad = (Model(initial_model.module, g[0]), Model(initial_model.module, g[1])) if lrs is not None else \
        Model(initial_model.module, g[0])
    return meta_grad, losses.detach(), (outer_loss.detach(), {k: v.detach() for k, v in outer_aux.items()})


# ---------------------------------------------------------------------------------------------- entry point
def multi_task_grad_and_losses(maml_def, key, initial_model, inner_lrs=None, task_slice=None, grad_scale=None):
    """maml.py:210-233 -> (grads, losses [T, K+1], (meta_losses [T], {term: [T]})).

    ``task_slice=(start, stop)`` evaluates only those tasks of the batch (task sharding across ranks,
    SURVEY.md 8(e)); ``grad_scale`` defaults to 1/n_batch_tasks, so the sum over ranks is the reference's mean."""
    T_all, K = maml_def.n_batch_tasks, maml_def.inner_steps
    keys = jr.split(key, T_all)
    lo, hi = task_slice if task_slice is not None else (0, T_all)
    scale = (1.0 / T_all) if grad_scale is None else grad_scale
    tasks = []
    for t in range(lo, hi):
        task_key, rollout_key, outer_loss_key = jr.split(keys[t], 3)
        tasks.append((maml_def.make_task_loss_fns(task_key), rollout_key, outer_loss_key))
    fns = [f for (pair, _, _) in tasks for f in pair]
    if _fusable(maml_def, fns):
        return _fused_outer(maml_def, tasks, initial_model, inner_lrs, K, scale)
    # generic route: one task at a time
    outs = [single_task_grad_and_losses(maml_def, keys[t], initial_model, inner_lrs) for t in range(lo, hi)]
    if inner_lrs is not None:
        grads = (Model(initial_model.module, scale * torch.stack([o[0][0].theta for o in outs]).sum(0)),
                 Model(initial_model.module, scale * torch.stack([o[0][1].theta for o in outs]).sum(0)))
    else:
        grads = Model(initial_model.module, scale * torch.stack([o[0].theta for o in outs]).sum(0))
    losses = torch.stack([o[1] for o in outs])
    meta = torch.stack([o[2][0] for o in outs])
    aux = {k: torch.stack([o[2][1][k] for o in outs]) for k in outs[0][2][1]}
    return grads, losses, (meta, aux)


def _gpu_sampled_sets(pde, tasks, K, K_):
    """All (2K + 2) x T point draws of one outer step on the GPU (compat/gpu_samplers.py; --gpu_samplers): the same keys the
    host route hands to ``sample`` (maml.py:132-140,158,192), one launch per distinct point count.  None when this PDE has
    no GPU sampler."""
    import numpy as np
    from . import gpu_samplers
    T = len(tasks)
    inner_keys, outer_keys = np.zeros((K + 1, T, 2), np.uint32), np.zeros((K + 1, T, 2), np.uint32)
    for t, ((inner_fn, outer_fn), rollout_key, outer_loss_key) in enumerate(tasks):
        k1s, k2s, final_key = _rollout_keys(rollout_key, K)
        for k in range(K):
            inner_keys[k, t], outer_keys[k, t] = k1s[k], k2s[k]
        inner_keys[K, t], outer_keys[K, t] = final_key, outer_loss_key
    (inner_fn0, outer_fn0) = tasks[0][0]
    params = [fn[0].params for fn, _, _ in tasks] * (K + 1)
    a = gpu_samplers.sample_points_batched(pde, inner_keys.reshape(-1, 2), inner_fn0.n_points, params, K_)
    if a is None:
        return None
    b = gpu_samplers.sample_points_batched(pde, outer_keys.reshape(-1, 2), outer_fn0.n_points, params, K_)
    split = lambda sets, k: [x.reshape(K + 1, T, *x.shape[1:])[k].contiguous() for x in sets]
    return ([split(a, k) for k in range(K)], [split(b, k) for k in range(K)], split(a, K), split(b, K))


def _fused_outer(maml_def, tasks, initial_model, inner_lrs, K, scale):
    dev = initial_model.theta.device
    K_ = ops.kernels(dev)
    inner_fn0, outer_fn0 = tasks[0][0]
    pde = inner_fn0.pde
    T = len(tasks)
    prms = [pde.abi_params(inner_fn.params) for (inner_fn, _), _, _ in tasks]
    sets = _gpu_sampled_sets(pde, tasks, K, K_) if FLAGS.gpu_samplers else None
    if sets is None:
        inner_pts, outer_pts = [[] for _ in range(K)], [[] for _ in range(K)]
        final_pts, aux_pts = [], []
        for (inner_fn, outer_fn), rollout_key, outer_loss_key in tasks:
            k1s, k2s, final_key = _rollout_keys(rollout_key, K)
            for k in range(K):
                inner_pts[k].append(inner_fn.sample(k1s[k]))
                outer_pts[k].append(outer_fn.sample(k2s[k]))
            final_pts.append(inner_fn.sample(final_key))
            aux_pts.append(outer_fn.sample(outer_loss_key))
        sets = ([_stack_points(pde, p) for p in inner_pts], [_stack_points(pde, p) for p in outer_pts],
                _stack_points(pde, final_pts), _stack_points(pde, aux_pts))
    inner_sets, outer_sets, final_sets, aux_sets = sets
    prm = torch.stack(prms).to(dev).contiguous()
    lrs = inner_lrs.theta.contiguous() if inner_lrs is not None else None
    # one packed buffer [g_theta | g_lrs]: the layout of the all-reduce and of the fused outer step (mpde_outer_step)
    P = initial_model.theta.numel()
    packed = torch.empty((1 + (K if lrs is not None else 0)) * P, dtype=torch.float32, device=dev)
    out = K_.maml_outer(initial_model.net_desc(K_), pde.name, inner_fn0.bc_weight, _cfg(maml_def, K),
                        initial_model.theta.detach().contiguous(), lrs, inner_sets, outer_sets, final_sets, aux_sets,
                        prm, T, grad_scale=scale, out_g_theta=packed[:P],
                        out_g_lrs=packed[P:].view(K, P) if lrs is not None else None)
    g_model = Model(initial_model.module, out["g_theta"])
    g_model.packed = packed
    grads = (g_model, Model(initial_model.module, out["g_lrs"])) if inner_lrs is not None else g_model
    aux = {n: out["aux_terms"][:, i] for i, n in enumerate(PDE_TERMS[pde.name])}
    return grads, out["losses"], (out["meta_losses"], aux)
