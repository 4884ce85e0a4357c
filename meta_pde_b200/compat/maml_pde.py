"""``python -m meta_pde_b200.compat.maml_pde`` -- drop-in for ``python -m src.maml_pde`` (src/maml_pde.py) on the
hot path: same flags, same MamlDef / Field construction, same train_step semantics (src/maml_pde.py:203-226), NaN stop,
text log and pickle checkpoints, and the deployment rollout of the n_eval validation tasks (get_final_model /
make_coef_func / validation_losses, src/maml_pde.py:163-201) with its ``deployment_time``.  The FEniCS ground truth the
reference compares those fields with, and the plots, are out of scope: ``val_mse`` etc. are logged
only when ground-truth values are supplied to ``trainer_util.vmap_validation_error`` by the caller."""
import os

import torch
from absl import app

from . import checkpoint, deploy, field, flags as cflags, maml, random as jr, trainer_util
from .flags import FLAGS
from .optim import Adam, Momentum
from .pdes import get_pde
from .tasks import make_task_loss_fns_factory
from .. import ops

cflags.define_maml()


def build(pde):
    """-> (maml_def, Field, optimizer, inner_lrs Model, lr-Adam state)  (src/maml_pde.py:100-159)"""
    maml_def = maml.MamlDef(make_inner_opt=Momentum(learning_rate=FLAGS.inner_lr, beta=0.0).create,
                            make_task_loss_fns=make_task_loss_fns_factory(pde), inner_steps=FLAGS.inner_steps,
                            n_batch_tasks=FLAGS.bsize, softplus_lrs=True, outer_loss_decay=FLAGS.outer_loss_decay)
    Field = pde.BaseField.partial(sizes=[FLAGS.layer_size for _ in range(FLAGS.num_layers)], dense_args=(),
                                  nonlinearity=field.sin if FLAGS.siren else field.swish, omega=FLAGS.siren_omega,
                                  omega0=FLAGS.siren_omega0, log_scale=FLAGS.log_scale)
    key, subkey = jr.split(jr.PRNGKey(0))
    _, init_params = Field.init_by_shape(subkey, [((1, 2), "float32")])
    model = field.Model(Field, init_params)
    inner_lrs = field.Model(Field, torch.ones(FLAGS.inner_steps, model.theta.numel(), device=model.theta.device))
    if FLAGS.load_model_from_expt is not None:
        model, lrs = checkpoint.load(FLAGS.load_model_from_expt, Field)
        inner_lrs = lrs if lrs is not None else inner_lrs
    optimizer = trainer_util.get_optimizer(Field, model)
    lr_state = {"m": torch.zeros_like(inner_lrs.theta), "v": torch.zeros_like(inner_lrs.theta)}
    return key, maml_def, Field, optimizer, inner_lrs, lr_state


def train_step(step, key, maml_def, optimizer, inner_lrs, lr_state):
    """src/maml_pde.py:203-226.  Returns (optimizer, inner_lrs, losses, meta_losses, meta_grad_norm)."""
    K = ops.kernels(optimizer.target.theta.device)
    (g_model, g_lrs), losses, meta_losses = maml.multi_task_grad_and_losses(maml_def, key, optimizer.target, inner_lrs)
    packed = getattr(g_model, "packed", None)            # fused route: [g_theta | g_lrs] already contiguous
    if packed is None:
        packed = torch.cat([g_model.theta.reshape(-1), g_lrs.theta.reshape(-1)]).contiguous()
    # One launch (mpde_outer_step): the norm of the MODEL grads only (quirk 1) rescales the lr-tree grads too, the outer
    # optimizer steps theta, jax.experimental.optimizers.adam at index `step` steps the lr tree (maml_pde.py:210-225)
    x = optimizer.target.theta.detach().clone().contiguous()
    lrs = inner_lrs.theta.detach().clone().contiguous()
    od = optimizer.optimizer_def
    t = optimizer.state["step"] + 1
    if isinstance(od, Adam):
        m, v = optimizer.state["m"].clone(), optimizer.state["v"].clone()
        sumsq = K.outer_step(x, packed, m, v, t, od.learning_rate, od.beta1, od.beta2, od.eps, FLAGS.grad_clip, "adam",
                             lrs=lrs.view(-1), m_lrs=lr_state["m"].view(-1), v_lrs=lr_state["v"].view(-1), t_lrs=step + 1,
                             lr_lrs=FLAGS.lr_inner_lr)
        state = {"step": t, "m": m, "v": v}
    else:
        sumsq = K.outer_step(x, packed, None, None, t, od.learning_rate, 0.0, 0.0, 1e-8, FLAGS.grad_clip, "sgd",
                             lrs=lrs.view(-1), m_lrs=lr_state["m"].view(-1), v_lrs=lr_state["v"].view(-1), t_lrs=step + 1,
                             lr_lrs=FLAGS.lr_inner_lr)
        state = {"step": t}
    optimizer = type(optimizer)(od, field.Model(optimizer.target.module, x), state)
    return optimizer, field.Model(inner_lrs.module, lrs), losses, meta_losses, torch.sqrt(sumsq)


def main(argv):
    if FLAGS.out_dir is None:
        FLAGS.out_dir = FLAGS.pde + "_maml_results"
    pde = get_pde(FLAGS.pde)
    path, log, _ = trainer_util.prepare_logging(FLAGS.out_dir, FLAGS.expt_name)
    log(FLAGS.flags_into_string())
    with open(os.path.join(path, "flags_config.txt"), "w") as f:
        f.write(FLAGS.flags_into_string())
    key, maml_def, Field, optimizer, inner_lrs, lr_state = build(pde)
    ev = deploy.make_maml_eval_fns(pde, maml_def)
    # validation tasks and their evaluation coordinates (maml_pde.py:228-236; coords come from FEniCS meshes in the
    # reference, from sample_points_in_domain here)
    key, gt_key, gt_points_key = jr.split(key, 3)
    gt_params = [pde.sample_params(k) for k in jr.split(gt_key, FLAGS.n_eval)]
    coords = torch.stack([torch.as_tensor(pde.sample_points_in_domain(k, FLAGS.validation_points, p), dtype=torch.float32)
                          for k, p in zip(jr.split(gt_points_key, FLAGS.n_eval), gt_params)])
    log("NN model:", {k: (tuple(v.shape) if torch.is_tensor(v) else {kk: tuple(vv.shape) for kk, vv in v.items()})
                      for k, v in optimizer.target.params.items()})
    for step in range(FLAGS.outer_steps):
        key, subkey = jr.split(key, 2)
        with trainer_util.Timer() as t:
            optimizer, inner_lrs, losses, meta_losses, meta_grad_norm = train_step(step, subkey, maml_def, optimizer,
                                                                                    inner_lrs, lr_state)
        if FLAGS.measure_grad_norm_every > 0 and step % FLAGS.measure_grad_norm_every == 0:
            log("loss vals and grad norms: ", {k: float(v.sum()) for k, v in meta_losses[1].items()})
            log("meta grad norm: ", float(meta_grad_norm))
        if bool(torch.isnan(meta_losses[0].mean())):
            log("encountered nan at at step {}".format(step))
            break
        if step % FLAGS.log_every == 0:
            with trainer_util.Timer() as deploy_timer:
                trainer_util.vmap_validation_rollout((optimizer.target, inner_lrs), gt_params, coords, None,
                                                     ev.partial_make_coef_func)
            deployment_time = deploy_timer.interval / FLAGS.n_eval
            val_losses, val_meta_losses = ev.validation_losses((optimizer.target, inner_lrs))
            log("step: {}, meta_loss: {}, val_meta_loss: {}, losses: {}, deployment_time: {}, meta_grad_norm: {}, "
                "time: {}".format(step, float(meta_losses[0].mean()), float(val_meta_losses[0].mean()),
                                  losses.mean(dim=0).tolist(), deployment_time, float(meta_grad_norm), t.interval))
        if FLAGS.viz_every > 0 and step % FLAGS.viz_every == 0:
            checkpoint.save(path, step, optimizer.target, inner_lrs)
    return optimizer, inner_lrs


if __name__ == "__main__":
    app.run(main)
