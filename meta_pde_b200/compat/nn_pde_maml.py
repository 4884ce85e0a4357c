"""``python -m meta_pde_b200.compat.nn_pde_maml`` -- drop-in for the README command ``python -m src.nn_pde_maml``
(src/nn_pde_maml.py): deploy a (meta-)initialisation on ONE PDE (n_eval = fixed_num_pdes = 1, :52-53), 5 hard-coded
first-order MAML steps (:178-241: Momentum 1e-4, K = 5, lr tree of ones or loaded), then Adam fine-tuning with the
batch loss SUMMED over bsize tasks (:110-124,140-171)."""
import os

import torch
from absl import app

from . import checkpoint, field, flags as cflags, maml, random as jr, trainer_util
from .flags import FLAGS
from .optim import Adam, Momentum
from .pdes import get_pde
from .tasks import PdeTaskLoss
from .. import ops
from .._abi import PDE_TERMS

cflags.define_nn()


def batch_loss_and_grad(pde, key, model):
    """value_and_grad(batch_loss_fn) (src/nn_pde_maml.py:110-124,142-144): one fused launch over the bsize tasks."""
    K = ops.kernels(model.theta.device)
    keys = jr.split(key, FLAGS.bsize)
    pts, prms = [], []
    for k in keys:
        k1, k2 = jr.split(k, 2)                            # task_loss_fn, :73-78
        params = pde.sample_params(k1)
        pts.append(pde.sample_points(k2, FLAGS.outer_points, params))
        prms.append(pde.abi_params(params))
    sets = maml._stack_points(pde, pts)
    prm = torch.stack(prms).to(model.theta.device).contiguous()
    loss, terms, grad = K.loss_grad(model.net_desc(K), pde.name, FLAGS.bc_weight, model.theta.contiguous(), sets, prm, FLAGS.bsize)
    aux = {n: terms[:, i].mean() for i, n in enumerate(PDE_TERMS[pde.name])}
    return loss.sum(), aux, K.task_sum(grad, 1.0)          # grad of the SUM over the bsize tasks (:110-124)


def train_step(pde, key, optimizer):
    """src/nn_pde_maml.py:140-171 -> (optimizer, loss, loss_aux, grad_norm)"""
    K = ops.kernels(optimizer.target.theta.device)
    loss, aux, grad = batch_loss_and_grad(pde, key, optimizer.target)
    od = optimizer.optimizer_def
    x = optimizer.target.theta.detach().clone().contiguous()
    g = grad.detach().contiguous().view(-1)
    t = optimizer.state["step"] + 1
    if isinstance(od, Adam):          # fused ||g||, clip, Adam (mpde_outer_step with no lr tree)
        m, v = optimizer.state["m"].clone(), optimizer.state["v"].clone()
        sumsq = K.outer_step(x, g, m, v, t, od.learning_rate, od.beta1, od.beta2, od.eps, FLAGS.grad_clip, "adam")
        state = {"step": t, "m": m, "v": v}
    else:
        sumsq = K.outer_step(x, g, None, None, t, od.learning_rate, 0.0, 0.0, 1e-8, FLAGS.grad_clip, "sgd")
        state = {"step": t}
    return type(optimizer)(od, field.Model(optimizer.target.module, x), state), loss, aux, torch.sqrt(sumsq)


def maml_step(pde, key, optimizer, inner_lrs=None):
    """src/nn_pde_maml.py:178-241: hard-coded MamlDef (ignores --inner_lr / --inner_steps: quirk 6)."""
    k1, _ = jr.split(key)
    kk1, _ = jr.split(k1, 2)                        # task_loss_fn(key, model) splits its key itself (:73-78)
    maml_def = maml.MamlDef(make_inner_opt=Momentum(learning_rate=1.0e-4, beta=0.0).create, make_task_loss_fns=None,
                            inner_steps=5, n_batch_tasks=1, softplus_lrs=True, outer_loss_decay=0.1)
    model = optimizer.target
    if inner_lrs is None:
        inner_lrs = field.Model(model.module, torch.ones(5, model.theta.numel(), device=model.theta.device))

    # fixed_num_pdes = 1 makes sample_params key-independent, so the fused adapt route applies with one params draw
    params = pde.sample_params(kk1)
    loss_fn = PdeTaskLoss(pde, params, FLAGS.outer_points)
    orig_sample = loss_fn.sample
    loss_fn.sample = lambda key: orig_sample(jr.split(key, 2)[1])
    final_model, (_, losses) = maml.single_task_rollout(maml_def, k1, model, loss_fn, inner_lrs)
    return trainer_util.get_optimizer(model.module, final_model), losses


def main(argv):
    FLAGS.n_eval = 1
    FLAGS.fixed_num_pdes = 1
    if FLAGS.out_dir is None:
        FLAGS.out_dir = FLAGS.pde + "_nn_results"
    pde = get_pde(FLAGS.pde)
    path, log, _ = trainer_util.prepare_logging(FLAGS.out_dir, FLAGS.expt_name)
    log(FLAGS.flags_into_string())
    Field = pde.BaseField.partial(sizes=[FLAGS.layer_size for _ in range(FLAGS.num_layers)], dense_args=(),
                                  nonlinearity=field.sin if FLAGS.siren else field.swish, omega=FLAGS.siren_omega,
                                  omega0=FLAGS.siren_omega0, log_scale=FLAGS.log_scale)
    key, subkey = jr.split(jr.PRNGKey(0))
    _, init_params = Field.init_by_shape(subkey, [((1, 2), "float32")])
    model, lrs = field.Model(Field, init_params), None
    if FLAGS.load_model_from_expt is not None:
        model, lrs = checkpoint.load(FLAGS.load_model_from_expt, Field)
    optimizer = trainer_util.get_optimizer(Field, model)
    key, _, _ = jr.split(key, 3)
    key, subkey = jr.split(key)
    optimizer, losses = maml_step(pde, subkey, optimizer, lrs)
    for i, l in enumerate(losses.tolist()):
        log("step: {}, loss: {}".format(i, l))
    optimizer, loss, loss_aux, grad_norm = train_step(pde, subkey, optimizer)      # reuses subkey (quirk 6)
    for step in range(5, FLAGS.outer_steps + 1):
        key, subkey = jr.split(key)
        with trainer_util.Timer() as t:
            optimizer, loss, loss_aux, grad_norm = train_step(pde, subkey, optimizer)
        if step % FLAGS.log_every == 0:
            log("step: {}, loss: {}, grad_norm: {}, per train step time: {}, {}".format(
                step, float(loss), float(grad_norm), t.interval, {k: float(v) for k, v in loss_aux.items()}))
        if FLAGS.viz_every > 0 and step % FLAGS.viz_every == 0:
            checkpoint.save(path, step, optimizer.target)
    return optimizer


if __name__ == "__main__":
    app.run(main)
