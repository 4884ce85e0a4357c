"""Oracle: MAML / LEAP rollouts and the drivers' train steps (TEST INFRASTRUCTURE ONLY).

Restates, with plain torch autograd through the unrolled inner loop:
  * maml_inner_step / single_task_rollout / single_task_grad_and_losses /
    multi_task_grad_and_losses           src/nets/maml.py:50-233
  * leap_inner_step / single_task_rollout / get_meta_grad_increment /
    compute_global_norm / multi_task_grad_and_losses   src/nets/leap.py:49-228
  * MAML train_step (clip by the MODEL-grad norm, Adam on theta, jax-adam on
    the lr tree)                          src/maml_pde.py:203-226
  * LEAP train_step                       src/leap_pde.py:175-195
  * fine-tune train_step / maml_step      src/nn_pde_maml.py:140-171,178-241

PRNG keys are replaced by explicit, pre-sampled point sets: the reference draws
fresh points from a key at every loss call; here the caller passes them in the
same order (sampler outputs are *inputs* to the oracle).

theta is the flat parameter vector of the C ABI (oracle/field.py); an "lr tree"
of the reference ([K, *leaf.shape] per leaf) is the flat [K, P] matrix.

Third-party optimiser formulas (flax 0.3.3 optim.Momentum/Adam, jax 0.2.20
experimental.optimizers.adam) are restated from those libraries' published
update rules; neither library is present in this image.
"""
from collections import namedtuple

import torch
import torch.nn.functional as F

from .field import unflatten_params, make_field_fn
from .pdes import driver_loss, PDE_OUT_DIM, PDE_TERMS

NetCfg = namedtuple("NetCfg", ["sizes", "omega", "log_scale", "out_dim", "nonlinearity"], defaults=("sin",))


def make_loss_fn(pde_name, net, bc_weight):
    """(theta, points, pde_params) -> (loss, aux dict): the closure of maml_pde.py:76-98."""

    def loss_fn(theta, points, pde_params):
        params = unflatten_params(theta, net.sizes, 2, net.out_dim, net.log_scale)
        field_fn = make_field_fn(params, net.out_dim, net.omega, net.nonlinearity)
        return driver_loss(pde_name, field_fn, points, pde_params, bc_weight)

    return loss_fn


# =============================================================== MAML
MamlCfg = namedtuple(
    "MamlCfg",
    ["inner_steps", "inner_lr", "inner_grad_clip", "softplus_lrs", "outer_loss_decay"],
)

# One task's pre-sampled data.  inner_points[k], outer_points[k] are the point
# tuples drawn from k1 / k2 of scan step k (maml.py:134-139); final_points from
# final_key (maml.py:158); aux_points from outer_loss_key (maml.py:192).
MamlTask = namedtuple("MamlTask", ["pde_params", "inner_points", "outer_points", "final_points", "aux_points"])


def global_norm(g):
    return torch.sqrt(torch.sum(g ** 2))


def maml_inner_step(theta, loss_of_theta, inner_lr_k, cfg, create_graph):
    """maml.py:50-97.  inner_lr_k: [P] tensor (lr tree row) or python float (scalar lr)."""
    loss, _ = loss_of_theta(theta)
    (grad,) = torch.autograd.grad(loss, theta, create_graph=create_graph)
    maybe_softplus = (lambda x: F.softplus(x)) if cfg.softplus_lrs else (lambda x: x)
    if torch.is_tensor(inner_lr_k):
        grad = grad * maybe_softplus(inner_lr_k)
    else:
        grad = grad * maybe_softplus(torch.tensor(float(inner_lr_k), dtype=theta.dtype))
    grad_norm = global_norm(grad)
    if bool(grad_norm > cfg.inner_grad_clip):  # lax.cond, maml.py:87-94
        grad = cfg.inner_grad_clip * grad / grad_norm
    new_theta = theta - cfg.inner_lr * grad  # flax Momentum(beta=0).apply_gradient
    return new_theta, loss, grad_norm


def maml_single_task_rollout(cfg, loss_fn, theta0, task, inner_lrs=None, with_outer=True, create_graph=True):
    """maml.py:100-166.  Returns theta_K, (meta_loss_sum, losses [K+1]), thetas [K+1]."""
    K = cfg.inner_steps
    theta = theta0
    meta_loss = torch.zeros((), dtype=theta0.dtype)
    losses, thetas = [], [theta0]
    for k in range(K):
        lr_k = inner_lrs[k] if inner_lrs is not None else 1.0  # np.ones(inner_steps), maml.py:126-127
        theta, loss, _ = maml_inner_step(
            theta, lambda th: loss_fn(th, task.inner_points[k], task.pde_params), lr_k, cfg, create_graph
        )
        losses.append(loss)
        thetas.append(theta)
        if with_outer:
            meta_loss = loss_fn(theta, task.outer_points[k], task.pde_params)[0] + meta_loss * cfg.outer_loss_decay
    loss_final, _ = loss_fn(theta, task.final_points, task.pde_params)
    losses.append(loss_final)
    return theta, (meta_loss, torch.stack(losses)), thetas


def maml_single_task_grad_and_losses(cfg, loss_fn, theta0, task, inner_lrs):
    """maml.py:169-207: grad of the rollout's meta-loss wrt (theta0, inner_lrs)."""
    theta0 = theta0.detach().clone().requires_grad_(True)
    lrs = inner_lrs.detach().clone().requires_grad_(True)
    theta_K, (outer_loss, losses), thetas = maml_single_task_rollout(cfg, loss_fn, theta0, task, lrs)
    _, outer_aux = loss_fn(theta_K, task.aux_points, task.pde_params)
    g_theta, g_lrs = torch.autograd.grad(outer_loss, (theta0, lrs), allow_unused=True)
    if g_lrs is None:
        g_lrs = torch.zeros_like(lrs)
    outer_aux = {k: v.detach() for k, v in outer_aux.items()}
    return (g_theta, g_lrs), losses.detach(), (outer_loss.detach(), outer_aux), [t.detach() for t in thetas]


def maml_multi_task_grad_and_losses(cfg, loss_fn, theta0, tasks, inner_lrs):
    """maml.py:210-233: vmap over tasks, mean of the grads over tasks."""
    outs = [maml_single_task_grad_and_losses(cfg, loss_fn, theta0, t, inner_lrs) for t in tasks]
    g_theta = torch.stack([o[0][0] for o in outs]).mean(dim=0)
    g_lrs = torch.stack([o[0][1] for o in outs]).mean(dim=0)
    losses = torch.stack([o[1] for o in outs])
    meta_losses = torch.stack([o[2][0] for o in outs])
    aux = {k: torch.stack([o[2][1][k] for o in outs]) for k in outs[0][2][1]}
    return (g_theta, g_lrs), losses, (meta_losses, aux)


# =============================================================== optimisers
class FlaxAdam:
    """flax 0.3.3 optim.Adam.apply_param_gradient (beta1, beta2, eps=1e-8, no weight decay)."""

    def __init__(self, lr, beta1=0.9, beta2=0.999, eps=1e-8):
        self.lr, self.b1, self.b2, self.eps = lr, beta1, beta2, eps

    def init(self, theta):
        return {"step": 0, "m": torch.zeros_like(theta), "v": torch.zeros_like(theta)}

    def apply(self, theta, grad, state):
        m = self.b1 * state["m"] + (1.0 - self.b1) * grad
        v = self.b2 * state["v"] + (1.0 - self.b2) * grad ** 2
        t = state["step"] + 1
        m_hat = m / (1.0 - self.b1 ** t)
        v_hat = v / (1.0 - self.b2 ** t)
        new_theta = theta - self.lr * m_hat / (torch.sqrt(v_hat) + self.eps)
        return new_theta, {"step": t, "m": m, "v": v}


class JaxAdam(FlaxAdam):
    """jax 0.2.20 experimental.optimizers.adam(step_size, b1, b2, eps=1e-8): same update rule,
    step index i passed by the caller (maml_pde.py:225)."""

    def apply_at(self, i, x, grad, state):
        st = dict(state)
        st["step"] = i
        return self.apply(x, grad, st)


def clip_by_norm(tree_vals, norm, clip):
    if bool(norm > clip):
        return [clip * x / norm for x in tree_vals]
    return list(tree_vals)


def maml_train_step(step, cfg, loss_fn, theta, lrs, tasks, opt_theta, st_theta, opt_lrs, st_lrs, grad_clip):
    """src/maml_pde.py:203-226."""
    (g_theta, g_lrs), losses, meta_losses = maml_multi_task_grad_and_losses(cfg, loss_fn, theta, tasks, lrs)
    meta_grad_norm = global_norm(g_theta)  # model part only (quirk 1)
    g_theta, g_lrs = clip_by_norm([g_theta, g_lrs], meta_grad_norm, grad_clip)
    theta, st_theta = opt_theta.apply(theta, g_theta, st_theta)
    lrs, st_lrs = opt_lrs.apply_at(step, lrs, g_lrs, st_lrs)
    return theta, lrs, st_theta, st_lrs, losses, meta_losses, meta_grad_norm


# =============================================================== LEAP
LeapCfg = namedtuple(
    "LeapCfg",
    ["inner_steps", "inner_lr", "inner_grad_clip", "norm", "loss_in_distance", "stabilize", "inner_opt"],
)  # inner_opt: "adam" (leap_pde.py:87) or "sgd"

# loss0_points from loss0_key (leap.py:124-126); step_points[k] from k1, new_points[k] from k2 (leap.py:67,93)
LeapTask = namedtuple("LeapTask", ["pde_params", "loss0_points", "step_points", "new_points"])


def leap_meta_grad_increment(cfg, new_theta, theta, new_loss, loss, grad):
    """leap.py:190-228"""
    d_loss = new_loss - loss
    if cfg.stabilize:
        d_loss = -torch.abs(d_loss)
    if cfg.norm:
        sum_sq = torch.sum((new_theta - theta) ** 2)
        if cfg.loss_in_distance:
            sum_sq = sum_sq + d_loss ** 2
        norm = torch.sqrt(sum_sq)
    else:
        norm = 1.0
    inc = theta - new_theta
    if cfg.loss_in_distance:
        inc = inc - d_loss * grad
    return inc / norm


def leap_single_task_rollout(cfg, loss_fn, theta0, task):
    """leap.py:105-147 (+ leap_inner_step :49-102).  Returns theta_K, meta_grad, losses [K+1]."""
    theta = theta0.detach().clone()
    loss0, _ = loss_fn(theta, task.loss0_points, task.pde_params)
    opt = FlaxAdam(cfg.inner_lr, 0.9, 0.99) if cfg.inner_opt == "adam" else None
    state = opt.init(theta) if opt is not None else None
    accum = torch.zeros_like(theta)
    losses = [loss0.detach()]
    for k in range(cfg.inner_steps):
        th = theta.detach().clone().requires_grad_(True)
        loss, _ = loss_fn(th, task.step_points[k], task.pde_params)
        (grad,) = torch.autograd.grad(loss, th)
        loss = loss.detach()
        grad_norm = global_norm(grad)
        if bool(grad_norm > cfg.inner_grad_clip):
            grad = cfg.inner_grad_clip * grad / grad_norm
        if opt is not None:
            new_theta, state = opt.apply(theta, grad, state)
        else:
            new_theta = theta - cfg.inner_lr * grad
        new_loss, _ = loss_fn(new_theta, task.new_points[k], task.pde_params)
        new_loss = new_loss.detach()
        accum = accum + leap_meta_grad_increment(cfg, new_theta, theta, new_loss, loss, grad)
        theta = new_theta
        losses.append(new_loss)
    return theta, accum, torch.stack(losses)


def leap_multi_task_grad_and_losses(cfg, loss_fn, theta0, tasks):
    """leap.py:170-187"""
    outs = [leap_single_task_rollout(cfg, loss_fn, theta0, t) for t in tasks]
    grads = torch.stack([o[1] for o in outs]).mean(dim=0)
    losses = torch.stack([o[2] for o in outs])
    return grads, losses


def leap_train_step(cfg, loss_fn, theta, tasks, opt_theta, st_theta, grad_clip):
    """src/leap_pde.py:175-195"""
    meta_grad, losses = leap_multi_task_grad_and_losses(cfg, loss_fn, theta, tasks)
    meta_grad_norm = global_norm(meta_grad)
    (meta_grad,) = clip_by_norm([meta_grad], meta_grad_norm, grad_clip)
    theta, st_theta = opt_theta.apply(theta, meta_grad, st_theta)
    return theta, st_theta, losses, meta_grad_norm


# =============================================================== fine-tuning (nn_pde_maml)
def finetune_train_step(loss_fn, theta, batch, opt_theta, st_theta, grad_clip):
    """src/nn_pde_maml.py:110-124,140-171: grad of the SUM over the task batch, clip, Adam.
    batch: list of (points, pde_params)."""
    th = theta.detach().clone().requires_grad_(True)
    per_task = [loss_fn(th, pts, prm) for pts, prm in batch]
    loss = torch.sum(torch.stack([p[0] for p in per_task]))
    loss_aux = {k: torch.mean(torch.stack([p[1][k] for p in per_task])).detach() for k in per_task[0][1]}
    (grad,) = torch.autograd.grad(loss, th)
    grad_norm = global_norm(grad)
    (grad,) = clip_by_norm([grad], grad_norm, grad_clip)
    theta, st_theta = opt_theta.apply(theta.detach(), grad, st_theta)
    return theta, st_theta, loss.detach(), loss_aux, grad_norm
