"""``python -m meta_pde_b200.compat.leap_pde`` -- drop-in for ``python -m src.leap_pde`` (src/leap_pde.py) on the hot
path: LeapDef with an Adam inner optimizer (leap_pde.py:87-97), train_step (leap_pde.py:175-195)."""
import os

import torch
from absl import app

from . import checkpoint, deploy, field, flags as cflags, leap, random as jr, trainer_util
from .flags import FLAGS
from .optim import Adam
from .pdes import get_pde
from .tasks import make_task_loss_fn_factory
from .. import ops

cflags.define_leap()


def build(pde):
    leap_def = leap.LeapDef(make_inner_opt=Adam(learning_rate=FLAGS.inner_lr, beta1=0.9, beta2=0.99).create,
                            make_task_loss_fn=make_task_loss_fn_factory(pde), inner_steps=FLAGS.inner_steps,
                            n_batch_tasks=FLAGS.bsize, norm=True, loss_in_distance=True, stabilize=True)
    Field = pde.BaseField.partial(sizes=[FLAGS.layer_size for _ in range(FLAGS.num_layers)], dense_args=(),
                                  nonlinearity=field.sin if FLAGS.siren else field.swish, omega=FLAGS.siren_omega,
                                  omega0=FLAGS.siren_omega0, log_scale=FLAGS.log_scale)
    key, subkey = jr.split(jr.PRNGKey(0))
    _, init_params = Field.init_by_shape(subkey, [((1, 2), "float32")])
    model = field.Model(Field, init_params)
    if FLAGS.load_model_from_expt is not None:
        model, _ = checkpoint.load(FLAGS.load_model_from_expt, Field)
    return key, leap_def, Field, trainer_util.get_optimizer(Field, model)


def train_step(key, leap_def, optimizer):
    """src/leap_pde.py:175-195 -> (optimizer, losses, meta_grad_norm)"""
    K = ops.kernels(optimizer.target.theta.device)
    meta_grad, losses = leap.multi_task_grad_and_losses(leap_def, key, optimizer.target)
    od = optimizer.optimizer_def
    x = optimizer.target.theta.detach().clone().contiguous()
    g = meta_grad.theta.detach().contiguous().view(-1)
    t = optimizer.state["step"] + 1
    if isinstance(od, Adam):          # fused ||g||, clip, Adam (mpde_outer_step with no lr tree)
        m, v = optimizer.state["m"].clone(), optimizer.state["v"].clone()
        sumsq = K.outer_step(x, g, m, v, t, od.learning_rate, od.beta1, od.beta2, od.eps, FLAGS.grad_clip, "adam")
        state = {"step": t, "m": m, "v": v}
    else:
        sumsq = K.outer_step(x, g, None, None, t, od.learning_rate, 0.0, 0.0, 1e-8, FLAGS.grad_clip, "sgd")
        state = {"step": t}
    return type(optimizer)(od, field.Model(optimizer.target.module, x), state), losses, torch.sqrt(sumsq)


def main(argv):
    if FLAGS.out_dir is None:
        FLAGS.out_dir = FLAGS.pde + "_leap_results"
    pde = get_pde(FLAGS.pde)
    path, log, _ = trainer_util.prepare_logging(FLAGS.out_dir, FLAGS.expt_name)
    log(FLAGS.flags_into_string())
    with open(os.path.join(path, "flags_config.txt"), "w") as f:
        f.write(FLAGS.flags_into_string())
    key, leap_def, Field, optimizer = build(pde)
    ev = deploy.make_leap_eval_fns(pde, leap_def)       # src/leap_pde.py:140-173
    key, gt_key, gt_points_key = jr.split(key, 3)
    gt_params = [pde.sample_params(k) for k in jr.split(gt_key, FLAGS.n_eval)]
    coords = torch.stack([torch.as_tensor(pde.sample_points_in_domain(k, FLAGS.validation_points, p), dtype=torch.float32)
                          for k, p in zip(jr.split(gt_points_key, FLAGS.n_eval), gt_params)])
    for step in range(FLAGS.outer_steps):
        key, subkey = jr.split(key, 2)
        with trainer_util.Timer() as t:
            optimizer, losses, meta_grad_norm = train_step(subkey, leap_def, optimizer)
        if bool(torch.isnan(losses.mean())):
            log("encountered nan at at step {}".format(step))
            break
        if step % FLAGS.log_every == 0:
            with trainer_util.Timer() as deploy_timer:                                     # leap_pde.py:224-230
                trainer_util.vmap_validation_rollout(optimizer.target, gt_params, coords, None, ev.partial_make_coef_func)
            deployment_time = deploy_timer.interval / FLAGS.n_eval
            log("step: {}, losses: {}, deployment_time: {}, meta_grad_norm: {}, per train step time: {}".format(
                step, losses.mean(dim=0).tolist(), deployment_time, float(meta_grad_norm), t.interval))
        if FLAGS.viz_every > 0 and step % FLAGS.viz_every == 0:
            checkpoint.save(path, step, optimizer.target)
    return optimizer


if __name__ == "__main__":
    app.run(main)
