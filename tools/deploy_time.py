"""deployment_time of the plugin surface (src/maml_pde.py:264-271): adapt the n_eval validation tasks with the learned
lr tree (get_final_model) and evaluate the adapted fields at the validation coordinates (make_coef_func), through
meta_pde_b200.compat.deploy -- the batched route vmap_validation_error uses (one mpde_maml_adapt + one mpde_field_eval)
and the per-task route.  Prints one JSON line per configuration (seconds per validation task, as the reference logs it).

    python tools/deploy_time.py [poisson|td_burgers|hyper_elasticity ...]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from meta_pde_b200.compat import flags as cflags  # noqa: E402

cflags.define_maml()
cflags.ensure_parsed(["deploy_time"])
from meta_pde_b200.compat import deploy, field, maml, pdes, random as jr, tasks, trainer_util  # noqa: E402
from meta_pde_b200.compat.flags import FLAGS  # noqa: E402
from meta_pde_b200.compat.optim import Momentum  # noqa: E402


def run(name, W, L, K, n_eval, n_pts, n_val):
    FLAGS.pde, FLAGS.n_eval, FLAGS.inner_points, FLAGS.validation_points = name, n_eval, n_pts, n_val
    FLAGS.siren_omega = FLAGS.siren_omega0 = 30.0
    FLAGS.bc_weight, FLAGS.xmin, FLAGS.inner_grad_clip = 1.0, -1.0, 100.0
    if name == "hyper_elasticity":
        FLAGS.max_holes = 5
    pde = pdes.get_pde(name)
    Field = pde.BaseField.partial(sizes=[W] * L, dense_args=(), nonlinearity=field.sin, omega=30.0, omega0=30.0, log_scale=True)
    _, params = Field.init_by_shape(jr.PRNGKey(0), [((1, 2), "float32")])
    model = field.Model(Field, params)
    lrs = field.Model(Field, torch.ones(K, model.theta.numel(), device=model.theta.device))
    mdef = maml.MamlDef(Momentum(learning_rate=1e-4, beta=0.0).create, tasks.make_task_loss_fns_factory(pde), K, n_eval, True, 0.1)
    ev = deploy.make_maml_eval_fns(pde, mdef)
    gt_params = [pde.sample_params(k) for k in jr.split(jr.PRNGKey(1), n_eval)]
    coords = torch.stack([torch.as_tensor(pde.sample_points_in_domain(k, n_val, p), dtype=torch.float32)
                          for k, p in zip(jr.split(jr.PRNGKey(2), n_eval), gt_params)])
    out = {}
    for label, fn in (("batched", ev.partial_make_coef_func),
                      ("per_task", lambda k, m, p, c: ev.make_coef_func(k, m, p, c, K, mdef))):
        ts = []
        for rep in range(4):
            with trainer_util.Timer() as t:
                c = trainer_util.vmap_validation_rollout((model, lrs), gt_params, coords, None, fn)
            ts.append(t.interval / n_eval)
        out[label] = {"deployment_time_s_per_task": min(ts[1:]), "first_call_s_per_task": ts[0]}
        assert torch.isfinite(c).all()
    # device-only part of the batched route (points already sampled): the kernels' share of deployment_time
    keys = jr.split(jr.PRNGKey(0), n_eval)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    thf = ev.get_final_models(keys, (model, lrs), gt_params, K)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print(json.dumps({"what": "deployment_time", "pde": name, "siren": f"{L}x{W}", "inner_steps": K, "n_eval": n_eval,
                      "inner_points": n_pts, "validation_points": int(coords.shape[1]), **out,
                      "note": "host threefry sampling of the inner points is inside deployment_time, as in the reference "
                              "(sample_points is inside get_final_model); get_final_models_s_per_task includes it",
                      "get_final_models_s_per_task": (t1 - t0) / n_eval, "theta_final_shape": list(thf.shape)}), flush=True)


if __name__ == "__main__":
    names = sys.argv[1:] or ["poisson", "td_burgers", "hyper_elasticity"]
    for nm in names:
        run(nm, 64, 3, 5, 16, 1024, 1024)
