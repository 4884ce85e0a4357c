"""GPU tests of the plugin-surface mirror (meta_pde_b200/compat): loss_fn through the fused kernels with torch
autograd to second order, the fused MAML / LEAP routes against the generic (autograd-orchestrated) routes, and the
three drivers end to end with the reference's flag strings."""
import pytest
import torch

from helpers import rel_l2
from meta_pde_b200.compat import flags as cflags

cflags.define_maml()
cflags.ensure_parsed(["test"])
from meta_pde_b200.compat import field, leap, maml, pdes, random as jr, tasks  # noqa: E402
from meta_pde_b200.compat.flags import FLAGS  # noqa: E402
from meta_pde_b200.compat.optim import Adam, Momentum  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture()
def flags_small():
    saved = {k: FLAGS[k].value for k in ("siren_omega", "siren_omega0", "bc_weight", "xmin", "inner_points", "outer_points",
                                         "inner_grad_clip", "max_holes", "layer_size", "num_layers")}
    FLAGS.siren_omega, FLAGS.siren_omega0, FLAGS.bc_weight, FLAGS.xmin = 30.0, 30.0, 1.0, -1.0
    FLAGS.inner_points = FLAGS.outer_points = 96
    FLAGS.max_holes = 5
    yield
    for k, v in saved.items():
        setattr(FLAGS, k, v)


def _model(pde, W=64, L=3, seed=0):
    Field = pde.BaseField.partial(sizes=[W] * L, dense_args=(), nonlinearity=field.sin, omega=FLAGS.siren_omega,
                                  omega0=FLAGS.siren_omega0, log_scale=True)
    _, params = Field.init_by_shape(jr.PRNGKey(seed), [((1, 2), "float32")])
    return field.Model(Field, params)


@pytest.mark.parametrize("name", ["poisson", "td_burgers", "hyper_elasticity"])
def test_loss_fn_autograd_to_second_order(gpu_kernels, flags_small, name):
    K = gpu_kernels
    pde = pdes.get_pde(name)
    model = _model(pde)
    prm = pde.sample_params(jr.PRNGKey(1))
    pts = pde.sample_points(jr.PRNGKey(2), 128, prm)
    theta = model.theta.clone().requires_grad_(True)
    b, d = pde.loss_fn(field.Model(model.module, theta), pts, prm)
    assert list(b) == pdes._BOUNDARY_TERMS[name] and list(d) == pdes._DOMAIN_TERMS[name]
    loss = 3.0 * torch.stack(list(b.values())).sum() + torch.stack(list(d.values())).sum()
    (g,) = torch.autograd.grad(loss, theta, create_graph=True)
    v = torch.randn_like(theta)
    (hv,) = torch.autograd.grad((g * v).sum(), theta)
    sets = [s.reshape(1, -1, 2).contiguous() for s in pde.abi_sets(pts)]
    l2, _, g2, hv2 = K.loss_hvp(model.net_desc(K), name, 3.0, model.theta, v.detach(), sets,
                                pde.abi_params(prm).reshape(1, -1).cuda(), 1)
    assert rel_l2(loss.detach().cpu(), l2[0].cpu()) < 1e-6
    assert rel_l2(g.detach().cpu(), g2[0].cpu()) < 1e-6 and rel_l2(hv.cpu(), hv2[0].cpu()) < 1e-6
    # field_fn(x) is the plain forward
    u = model(pts[-1])
    assert u.shape[0] == pts[-1].shape[0] and torch.isfinite(u).all()
    with pytest.raises(TypeError):
        pde.loss_fn(lambda x: x.sum(-1), pts, prm)             # opaque callables cannot be fused: no fallback


def test_maml_fused_route_matches_generic_route(gpu_kernels, flags_small):
    pde = pdes.get_pde("poisson")
    model = _model(pde)
    T, Ksteps = 3, 2
    P = model.theta.numel()
    lrs = field.Model(model.module, 1.0 + 0.2 * torch.randn(Ksteps, P, device="cuda"))
    mk = tasks.make_task_loss_fns_factory(pde)
    mdef = maml.MamlDef(Momentum(learning_rate=1e-2, beta=0.0).create, mk, Ksteps, T, True, 0.1)
    # generic: hide the PdeTaskLoss type so that the autograd-orchestrated route runs
    hide = lambda key: tuple((lambda f: (lambda k, m: f(k, m)))(f) for f in mk(key))
    gdef = mdef._replace(make_task_loss_fns=hide)
    for clip in (1e14, 1e-2):
        FLAGS.inner_grad_clip = clip
        (gm, gl), losses, (meta, aux) = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs)
        (gm2, gl2), losses2, (meta2, aux2) = maml.multi_task_grad_and_losses(gdef, jr.PRNGKey(5), model, lrs)
        errs = dict(g_theta=rel_l2(gm.theta.cpu(), gm2.theta.cpu()), g_lrs=rel_l2(gl.theta.cpu(), gl2.theta.cpu()),
                    losses=rel_l2(losses.cpu(), losses2.cpu()), meta=rel_l2(meta.cpu(), meta2.cpu()),
                    aux=rel_l2(aux["domain_loss"].cpu(), aux2["domain_loss"].cpu()))
        assert all(e < 1e-4 for e in errs.values()), (clip, errs)
    # sharding: two halves with the global scale add up to the full batch
    FLAGS.inner_grad_clip = 1e14
    (ga, _), la, _ = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs, task_slice=(0, 2))
    (gb, _), lb, _ = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs, task_slice=(2, 3))
    (gm, _), losses, _ = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs)
    assert rel_l2((ga.theta + gb.theta).cpu(), gm.theta.cpu()) < 1e-5 and torch.equal(torch.cat([la, lb]), losses)
    # deployment rollout without lr tree (scalar softplus(1) lr) through both routes
    fm, (_, l1) = maml.single_task_rollout(mdef, jr.PRNGKey(6), model, mk(jr.PRNGKey(7))[0])
    fm2, (_, l2) = maml.single_task_rollout(gdef, jr.PRNGKey(6), model, hide(jr.PRNGKey(7))[0])
    assert rel_l2((fm.theta - model.theta).cpu(), (fm2.theta - model.theta).cpu()) < 1e-4 and rel_l2(l1.cpu(), l2.cpu()) < 1e-5


@pytest.mark.parametrize("opt", ["adam", "sgd"])
def test_leap_fused_route_matches_generic_route(gpu_kernels, flags_small, opt):
    pde = pdes.get_pde("td_burgers")
    FLAGS.inner_points = 126
    model = _model(pde)
    mk = tasks.make_task_loss_fn_factory(pde)
    od = Adam(learning_rate=1e-3, beta1=0.9, beta2=0.99) if opt == "adam" else Momentum(learning_rate=1e-2, beta=0.0)
    ldef = leap.LeapDef(od.create, mk, 3, 2, True, True, True)
    gdef = ldef._replace(make_task_loss_fn=lambda key: (lambda f: (lambda k, m: f(k, m)))(mk(key)))
    g1, l1 = leap.multi_task_grad_and_losses(ldef, jr.PRNGKey(8), model)
    g2, l2 = leap.multi_task_grad_and_losses(gdef, jr.PRNGKey(8), model)
    assert rel_l2(g1.theta.cpu(), g2.theta.cpu()) < 1e-4 and rel_l2(l1.cpu(), l2.cpu()) < 1e-5


def _run(module, argv, tmp_path):
    saved = {k: FLAGS[k].value for k in FLAGS}
    try:
        FLAGS(["x"] + argv + [f"--out_dir={tmp_path}", "--expt_name=t"])
        return module.main([])
    finally:
        for k, v in saved.items():
            try:
                setattr(FLAGS, k, v)
            except Exception:
                pass


def test_drivers_run_with_reference_flag_strings(gpu_kernels, tmp_path):
    from meta_pde_b200.compat import checkpoint, leap_pde, maml_pde, nn_pde_maml
    common = ["--pde=poisson", "--siren_omega=30.", "--siren_omega0=30.", "--bc_weight=1.0", "--optimizer=adam", "--xmin=-1.0",
              "--num_layers=3", "--layer_size=64", "--inner_points=128", "--outer_points=128", "--log_every=1", "--viz_every=2"]
    opt, lrs = _run(maml_pde, common + ["--bsize=4", "--inner_steps=2", "--outer_steps=3", "--inner_lr=1e-4", "--outer_lr=1e-4",
                                        "--inner_grad_clip=100", "--grad_clip=100"], tmp_path / "maml")
    assert torch.isfinite(opt.target.theta).all() and not torch.equal(lrs.theta, torch.ones_like(lrs.theta))
    m, l = checkpoint.load(str(tmp_path / "maml" / "t"), opt.target.module)
    assert l is not None and m.theta.shape == opt.target.theta.shape                # [model, lrs] pickle of step 2
    opt2 = _run(leap_pde, common + ["--bsize=2", "--inner_steps=3", "--outer_steps=2", "--inner_lr=1e-4"], tmp_path / "leap")
    assert torch.isfinite(opt2.target.theta).all()
    # README command shape: fine-tune from the MAML checkpoint (5 hard-coded MAML steps, then Adam)
    opt3 = _run(nn_pde_maml, common + ["--bsize=1", "--outer_steps=8", "--outer_lr=1e-5",
                                       f"--load_model_from_expt={tmp_path / 'maml' / 't'}"], tmp_path / "nn")
    assert torch.isfinite(opt3.target.theta).all()
    assert "step: 8" in open(tmp_path / "nn" / "t" / "log.txt").read()


def test_gpu_samplers_match_host_samplers(gpu_kernels, flags_small):
    """csrc/sampler_kernels.cuh through the C ABI: threefry bit-exact, Poisson / Burgers points = the host samplers' (which
    are pinned to the reference's sampler source), and the fused MAML route gives the same meta-gradient with
    --gpu_samplers as with the host samplers (same keys -> same points -> same numbers)."""
    import sampler_cases as sc
    sc.check_threefry(gpu_kernels)
    sc.check_poisson(gpu_kernels, n=256, S=8)
    sc.check_burgers(gpu_kernels)
    pde = pdes.get_pde("poisson")
    model = _model(pde)
    T, Ksteps = 3, 2
    lrs = field.Model(model.module, torch.ones(Ksteps, model.theta.numel(), device="cuda"))
    mdef = maml.MamlDef(Momentum(learning_rate=1e-2, beta=0.0).create, tasks.make_task_loss_fns_factory(pde), Ksteps, T, True, 0.1)
    saved = FLAGS.gpu_samplers
    try:
        FLAGS.gpu_samplers = False
        (g0, l0), losses0, (meta0, _) = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs)
        FLAGS.gpu_samplers = True
        (g1, l1), losses1, (meta1, _) = maml.multi_task_grad_and_losses(mdef, jr.PRNGKey(5), model, lrs)
    finally:
        FLAGS.gpu_samplers = saved
    # LEAP's fused route (td_burgers: bit-exact points -> identical numbers)
    bpde = pdes.get_pde("td_burgers")
    FLAGS.inner_points = 126
    bmodel = _model(bpde)
    ldef = leap.LeapDef(Adam(learning_rate=1e-3, beta1=0.9, beta2=0.99).create, tasks.make_task_loss_fn_factory(bpde), 3, 2, True, True, True)
    try:
        FLAGS.gpu_samplers = False
        ga, la = leap.multi_task_grad_and_losses(ldef, jr.PRNGKey(8), bmodel)
        FLAGS.gpu_samplers = True
        gb, lb = leap.multi_task_grad_and_losses(ldef, jr.PRNGKey(8), bmodel)
    finally:
        FLAGS.gpu_samplers = saved
    assert torch.equal(ga.theta, gb.theta) and torch.equal(la, lb)
    # boundary points differ by ~1e-7 (cos / sin), the drawn candidates are the same
    assert rel_l2(g1.theta.cpu(), g0.theta.cpu()) < 1e-4 and rel_l2(l1.theta.cpu(), l0.theta.cpu()) < 1e-4
    assert rel_l2(losses1.cpu(), losses0.cpu()) < 1e-5 and rel_l2(meta1.cpu(), meta0.cpu()) < 1e-5


@pytest.mark.parametrize("name", ["poisson", "td_burgers"])
def test_deployment_rollout_matches_oracle(gpu_kernels, flags_small, name):
    """get_final_model / make_coef_func / vmap_validation_error (src/maml_pde.py:163-201, trainer_util.py:501-575):
    the batched route (one mpde_maml_adapt + one mpde_field_eval for all n_eval tasks) and the per-task route against
    the float64 oracle rollout on the same sampled points, same learned lr tree."""
    from helpers import oracle_points, oracle_params
    from meta_pde_b200.compat import deploy
    from oracle.meta import MamlCfg, MamlTask, NetCfg, make_loss_fn, maml_single_task_rollout
    from oracle.field import make_field_fn, unflatten_params

    saved = {k: FLAGS[k].value for k in ("n_eval", "pde", "num_tsteps", "inner_lr")}
    try:
        FLAGS.n_eval, FLAGS.pde, FLAGS.num_tsteps = 3, name, 4
        if name == "td_burgers":
            FLAGS.inner_points = 126
        pde = pdes.get_pde(name)
        W, L, Ksteps = 64, 3, 3
        model = _model(pde, W, L)
        P = model.theta.numel()
        lrs = field.Model(model.module, 1.0 + 0.2 * torch.randn(Ksteps + 1, P, device="cuda"))
        mdef = maml.MamlDef(Momentum(learning_rate=1e-2, beta=0.0).create, tasks.make_task_loss_fns_factory(pde), Ksteps,
                            FLAGS.n_eval, True, 0.1)
        ev = deploy.make_maml_eval_fns(pde, mdef)
        gt_params = [pde.sample_params(k) for k in jr.split(jr.PRNGKey(11), FLAGS.n_eval)]
        coords = torch.stack([torch.as_tensor(pde.sample_points_in_domain(k, 64, p), dtype=torch.float32)
                              for k, p in zip(jr.split(jr.PRNGKey(12), FLAGS.n_eval), gt_params)])
        n_c = coords.shape[1]
        keys = jr.split(jr.PRNGKey(0), FLAGS.n_eval)          # trainer_util.py:505-506
        # ---- oracle: same key derivations, float64 rollout, plain forward ----
        net = NetCfg([W] * L, FLAGS.siren_omega, True, 1)
        lf = make_loss_fn(name, net, float(FLAGS.bc_weight))
        cfg = MamlCfg(Ksteps, 1e-2, float(FLAGS.inner_grad_clip), True, 0.1)
        want = []
        for i in range(FLAGS.n_eval):
            k1, _ = jr.split(keys[i], 2)
            pts = pde.sample_points(k1, FLAGS.inner_points, gt_params[i])
            sets = [torch.as_tensor(s, dtype=torch.float32).cpu().reshape(1, -1, 2) for s in pde.abi_sets(pts)]
            prm = pde.abi_params(gt_params[i]).cpu().reshape(1, -1)
            op, opr = oracle_points(name, sets, 0), oracle_params(name, prm, 0)
            task = MamlTask(opr, [op] * Ksteps, None, op, None)
            thK, _, _ = maml_single_task_rollout(cfg, lf, model.theta.double().cpu().requires_grad_(True), task, lrs.theta[:Ksteps].double().cpu(),
                                                 with_outer=False, create_graph=False)
            f = make_field_fn(unflatten_params(thK.detach(), [W] * L, 2, 1, True), 1, FLAGS.siren_omega)
            want.append(f(coords[i].double().cpu()).reshape(-1))
        want = torch.stack(want)
        # ---- batched route (what vmap_validation_error uses) and per-task route ----
        got_b = ev.make_coef_funcs(keys, (model, lrs), gt_params, coords).reshape(FLAGS.n_eval, n_c).cpu()
        got_s = torch.stack([ev.partial_make_coef_func(keys[i], (model, lrs), gt_params[i], coords[i].cuda()).reshape(-1)
                             for i in range(FLAGS.n_eval)]).cpu()
        assert rel_l2(got_b, want) < 1e-4 and rel_l2(got_s, want) < 1e-4, (rel_l2(got_b, want), rel_l2(got_s, want))
        # the adaptation moved the field (otherwise the comparison would be vacuous)
        base = model(coords[0].cuda()).reshape(-1).cpu()
        assert rel_l2(got_b[0], base) > 1e-5
        # ---- error arithmetic against a numpy restatement with synthetic ground truth ----
        import numpy as np
        gt_vals = (want + 0.1 * torch.randn_like(want)).numpy()
        mse, norms, rel, per_dim, rel_std, t_rel = deploy.vmap_validation_error((model, lrs), gt_params, coords, gt_vals,
                                                                                 ev.partial_make_coef_func)
        c = want.numpy().reshape(FLAGS.n_eval, n_c, 1); g = gt_vals.reshape(c.shape)
        err = c - g
        normalizer = (g ** 2).mean(axis=1, keepdims=True)
        rsq = err ** 2 / normalizer.mean(axis=2, keepdims=True)
        assert abs(float(mse) - (err ** 2).mean()) < 1e-4 * (err ** 2).mean()
        assert abs(float(rel) - rsq.mean()) < 1e-3 * rsq.mean()
        assert abs(float(rel_std) - rsq.mean(axis=(1, 2)).std()) < 1e-3 * max(rsq.mean(axis=(1, 2)).std(), 1e-12)
        assert (t_rel is not None) == (name == "td_burgers")
        if t_rel is not None:
            assert t_rel.shape == (FLAGS.num_tsteps,) and torch.isfinite(t_rel).all()
    finally:
        for k, v in saved.items():
            setattr(FLAGS, k, v)
