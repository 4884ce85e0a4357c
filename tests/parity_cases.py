"""Parity checkers shared by the emulator (CPU) and GPU test modules.

Each checker runs the kernels through the C ABI (``K`` = meta_pde_b200._abi.Kernels bound to
either the CUDA library or the test-only emulator build) and the float64 oracle on the same
seeded inputs, and asserts the north-star tolerance: relative 1e-4 (vector rel-L2) on
residuals, losses, inner-updated params and outer gradients.
"""
import torch

from helpers import oracle_params, oracle_points, rel_l2, synthetic
from meta_pde_b200._abi import PDE_OUT_DIM, PDE_TERMS
from oracle import meta as ometa
from oracle.field import make_field_fn, unflatten_params
from oracle import pdes as opdes

TOL = 1e-4  # BASELINE.json north_star: "relative 1e-4 on residuals, losses, inner-updated params and outer gradients"
OMEGA = 30.0
ACT = "sin"   # hidden nonlinearity of the cases below; `with nonlinearity("swish"):` runs them on the `--nosiren` network


class nonlinearity:
    """Run the checkers of this module with another hidden nonlinearity (maml_pde.py:115)."""

    def __init__(self, act):
        self.act = act

    def __enter__(self):
        global ACT
        self.prev, ACT = ACT, self.act

    def __exit__(self, *exc):
        global ACT
        ACT = self.prev


def _theta(sizes, out_dim, omega, omega0, log_scale, seed=0, bias_scale=0.0):
    if ACT == "sin":
        return synthetic.siren_theta(sizes, out_dim, omega, omega0, log_scale, seed=seed, bias_scale=bias_scale)
    return synthetic.swish_theta(sizes, out_dim, log_scale, seed=seed, bias_scale=bias_scale)


def _tol(name, theta_K, theta0):
    """1e-4, except for the UPDATE theta_K - theta_0, whose float32 resolution is eps32 * |theta| / |update|."""
    if name != "theta_delta":
        return TOL
    d = (theta_K.double() - theta0.double()).norm().item()
    return max(TOL, 8 * 6e-8 * theta_K.double().norm().item() / max(d, 1e-300))


def _dev(K, x):
    return x.to(K.device)


def make_problem(pde, W, L, n, T, seed=0, per_task_theta=True, set_sizes=None):
    out = PDE_OUT_DIM[pde]
    theta = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    if per_task_theta:
        g = torch.Generator().manual_seed(seed + 100)
        theta = torch.stack([theta * (1.0 + 0.02 * t) + 1e-3 * torch.randn(theta.numel(), generator=g) for t in range(T)])
    sizes = set_sizes if set_sizes is not None else synthetic.default_set_sizes(pde, n)
    sets, prm = synthetic.make_tasks(pde, T, sizes, seed=seed + 1)
    return theta, sets, prm


def oracle_loss_fn(pde, W, L, bc_weight=1.0):
    return ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, OMEGA, True, PDE_OUT_DIM[pde], ACT), bc_weight)


def check_loss_grad_hvp(K, pde, W, L, n, T, bc_weight=1.0, seed=0, set_sizes=None, per_task_theta=True):
    theta, sets, prm = make_problem(pde, W, L, n, T, seed, per_task_theta, set_sizes)
    P = theta.shape[-1]
    v = torch.randn(T, P, generator=torch.Generator().manual_seed(seed + 7))
    net = K.net_desc(L, W, PDE_OUT_DIM[pde], OMEGA, True, nonlinearity=ACT)
    dsets = [_dev(K, s) for s in sets]
    loss0, terms0, res = K.loss(net, pde, bc_weight, _dev(K, theta), dsets, _dev(K, prm), T, want_residual=True)
    loss1, terms1, grad1 = K.loss_grad(net, pde, bc_weight, _dev(K, theta), dsets, _dev(K, prm), T)
    loss2, terms2, grad2, hv = K.loss_hvp(net, pde, bc_weight, _dev(K, theta), _dev(K, v), dsets, _dev(K, prm), T)
    lf = oracle_loss_fn(pde, W, L, bc_weight)
    errs = {}
    for t in range(T):
        th0 = theta[t] if theta.dim() == 2 else theta
        th = th0.double().requires_grad_(True)
        pts, prm_t = oracle_points(pde, sets, t), oracle_params(pde, prm, t)
        l, aux = lf(th, pts, prm_t)
        (g,) = torch.autograd.grad(l, th, create_graph=True)
        (h,) = torch.autograd.grad((g * v[t].double()).sum(), th)
        ot = torch.stack([aux[k].detach() for k in PDE_TERMS[pde]])
        # per-point residual / energy density
        params = unflatten_params(th0.double(), [W] * L, 2, PDE_OUT_DIM[pde], True)
        ffn = make_field_fn(params, PDE_OUT_DIM[pde], OMEGA, ACT)
        if pde == "poisson":
            r = opdes.poisson_residual(ffn, pts[1], prm_t[0])
        elif pde == "td_burgers":
            r = opdes.burgers_rhs(ffn, pts[3], prm_t)
        else:
            r = opdes.hyper_energy_density(ffn, pts[5], prm_t)
        for name, got, want in [
            ("loss0", loss0[t], l), ("loss1", loss1[t], l), ("loss2", loss2[t], l),
            ("terms0", terms0[t], ot), ("terms1", terms1[t], ot), ("terms2", terms2[t], ot),
            ("residual", res[t], r), ("grad1", grad1[t], g), ("grad2", grad2[t], g), ("hv", hv[t], h),
        ]:
            e = rel_l2(got.detach().cpu(), want.detach())
            errs[name] = max(errs.get(name, 0.0), e)
    bad = {k: e for k, e in errs.items() if not (e < TOL)}
    assert not bad, f"{pde} W={W} L={L} n={n}: rel errors above {TOL}: {bad} (all: {errs})"
    return errs


def make_maml_data(pde, n_in, n_out, T, K_steps, seed):
    """Per-step inner/outer point sets + final/aux sets, ABI layout (each a list of [T, n, 2] tensors)."""
    si, so = synthetic.default_set_sizes(pde, n_in), synthetic.default_set_sizes(pde, n_out)
    inner, outer = [], []
    prm = None
    for k in range(K_steps):
        s, prm_k = synthetic.make_tasks(pde, T, si, seed=seed + 10 * k + 1)
        inner.append(s)
        prm = prm_k if prm is None else prm
        s2, _ = synthetic.make_tasks(pde, T, so, seed=seed + 10 * k + 2)
        outer.append(s2)
    final, _ = synthetic.make_tasks(pde, T, si, seed=seed + 1001)
    aux, _ = synthetic.make_tasks(pde, T, so, seed=seed + 1002)
    return inner, outer, final, aux, prm


def check_maml(K, pde, W, L, n, T, K_steps=2, inner_lr=1e-3, clip=1e14, decay=0.1, lr_tree=True, seed=0, bc_weight=1.0):
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    P = theta0.numel()
    inner, outer, final, aux, prm = make_maml_data(pde, n, n, T, K_steps, seed)
    g = torch.Generator().manual_seed(seed + 55)
    lrs = (1.0 + 0.3 * torch.randn(K_steps, P, generator=g)) if lr_tree else None
    cfg = dict(inner_steps=K_steps, inner_lr=inner_lr, inner_grad_clip=clip, softplus_lrs=True, outer_loss_decay=decay)
    net = K.net_desc(L, W, out, OMEGA, True, nonlinearity=ACT)
    d = lambda sets: [_dev(K, s) for s in sets]
    got = K.maml_outer(net, pde, bc_weight, cfg, _dev(K, theta0), _dev(K, lrs) if lrs is not None else None,
                       [d(s) for s in inner], [d(s) for s in outer], d(final), d(aux), _dev(K, prm), T,
                       want_theta_final=True)
    # oracle
    lf = oracle_loss_fn(pde, W, L, bc_weight)
    ocfg = ometa.MamlCfg(K_steps, inner_lr, clip, True, decay)
    tasks = []
    for t in range(T):
        tasks.append(ometa.MamlTask(oracle_params(pde, prm, t), [oracle_points(pde, s, t) for s in inner],
                                    [oracle_points(pde, s, t) for s in outer], oracle_points(pde, final, t),
                                    oracle_points(pde, aux, t)))
    olrs = lrs.double() if lrs is not None else torch.ones(K_steps, 1, dtype=torch.float64).expand(K_steps, P).contiguous()
    # scalar-lr variant of the reference uses np.ones(K): softplus(1.0) on every parameter == lr tree of ones
    outs = [ometa.maml_single_task_grad_and_losses(ocfg, lf, theta0.double(), tk, olrs) for tk in tasks]
    g_theta = torch.stack([o[0][0] for o in outs]).mean(0)
    g_lrs = torch.stack([o[0][1] for o in outs]).mean(0)
    losses = torch.stack([o[1] for o in outs])
    meta = torch.stack([o[2][0] for o in outs])
    auxd = torch.stack([torch.stack([o[2][1][k] for k in PDE_TERMS[pde]]) for o in outs])
    theta_K = torch.stack([o[3][-1] for o in outs])
    clipped = []
    errs = {
        "g_theta": rel_l2(got["g_theta"].cpu(), g_theta),
        "losses": rel_l2(got["losses"].cpu(), losses),
        "meta_losses": rel_l2(got["meta_losses"].cpu(), meta),
        "aux_terms": rel_l2(got["aux_terms"].cpu(), auxd),
        "theta_final": rel_l2(got["theta_final"].cpu(), theta_K),
        # inner-updated params: compare the UPDATE (theta_K - theta_0), the part the kernels produce
        "theta_delta": rel_l2(got["theta_final"].cpu().double() - theta0.double(), theta_K - theta0.double()),
    }
    if lrs is not None:
        errs["g_lrs"] = rel_l2(got["g_lrs"].cpu(), g_lrs)
    bad = {k: e for k, e in errs.items() if not (e < _tol(k, theta_K, theta0))}
    assert not bad, f"MAML {pde} W={W} L={L}: rel errors above {TOL}: {bad} (all: {errs})"
    return errs


def check_maml_adapt(K, pde, W, L, n, T, K_steps=3, inner_lr=1e-3, clip=1e14, seed=0):
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    inner, _, final, _, prm = make_maml_data(pde, n, n, T, K_steps, seed)
    cfg = dict(inner_steps=K_steps, inner_lr=inner_lr, inner_grad_clip=clip, softplus_lrs=True, outer_loss_decay=0.0)
    net = K.net_desc(L, W, out, OMEGA, True, nonlinearity=ACT)
    d = lambda sets: [_dev(K, s) for s in sets]
    thf, losses = K.maml_adapt(net, pde, 1.0, cfg, _dev(K, theta0), None, [d(s) for s in inner], d(final), _dev(K, prm), T)
    lf = oracle_loss_fn(pde, W, L, 1.0)
    ocfg = ometa.MamlCfg(K_steps, inner_lr, clip, True, 0.0)
    errs = {"theta": 0.0, "delta": 0.0, "losses": 0.0}
    for t in range(T):
        tk = ometa.MamlTask(oracle_params(pde, prm, t), [oracle_points(pde, s, t) for s in inner], None,
                            oracle_points(pde, final, t), None)
        th = theta0.double().requires_grad_(True)
        thK, (_, ls), _ = ometa.maml_single_task_rollout(ocfg, lf, th, tk, None, with_outer=False, create_graph=False)
        errs["theta"] = max(errs["theta"], rel_l2(thf[t].cpu(), thK.detach()))
        errs["delta"] = max(errs["delta"], rel_l2(thf[t].cpu().double() - theta0.double(), thK.detach() - theta0.double()))
        errs["losses"] = max(errs["losses"], rel_l2(losses[t].cpu(), ls.detach()))
    bad = {k: e for k, e in errs.items() if not (e < TOL)}
    assert not bad, f"MAML adapt {pde}: {bad} (all: {errs})"
    return errs


def check_leap(K, pde, W, L, n, T, K_steps=3, inner_lr=1e-3, clip=1e14, inner_opt="adam", seed=0):
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    sizes = synthetic.default_set_sizes(pde, n)
    loss0, prm = synthetic.make_tasks(pde, T, sizes, seed=seed + 1)
    step = [synthetic.make_tasks(pde, T, sizes, seed=seed + 20 + k)[0] for k in range(K_steps)]
    new = [synthetic.make_tasks(pde, T, sizes, seed=seed + 40 + k)[0] for k in range(K_steps)]
    cfg = dict(inner_steps=K_steps, inner_lr=inner_lr, inner_grad_clip=clip, norm=True, loss_in_distance=True,
               stabilize=True, inner_opt=inner_opt, beta1=0.9, beta2=0.99, eps=1e-8)
    net = K.net_desc(L, W, out, OMEGA, True, nonlinearity=ACT)
    d = lambda sets: [_dev(K, s) for s in sets]
    got = K.leap_outer(net, pde, 1.0, cfg, _dev(K, theta0), d(loss0), [d(s) for s in step], [d(s) for s in new],
                       _dev(K, prm), T, want_theta_final=True)
    lf = oracle_loss_fn(pde, W, L, 1.0)
    ocfg = ometa.LeapCfg(K_steps, inner_lr, clip, True, True, True, inner_opt)
    tasks = [ometa.LeapTask(oracle_params(pde, prm, t), oracle_points(pde, loss0, t),
                            [oracle_points(pde, s, t) for s in step], [oracle_points(pde, s, t) for s in new])
             for t in range(T)]
    outs = [ometa.leap_single_task_rollout(ocfg, lf, theta0.double(), tk) for tk in tasks]
    mg = torch.stack([o[1] for o in outs]).mean(0)
    losses = torch.stack([o[2] for o in outs])
    thK = torch.stack([o[0] for o in outs])
    errs = {
        "meta_grad": rel_l2(got["meta_grad"].cpu(), mg),
        "losses": rel_l2(got["losses"].cpu(), losses),
        "theta_final": rel_l2(got["theta_final"].cpu(), thK),
        "theta_delta": rel_l2(got["theta_final"].cpu().double() - theta0.double(), thK - theta0.double()),
    }
    bad = {k: e for k, e in errs.items() if not (e < _tol(k, thK, theta0))}
    assert not bad, f"LEAP {pde} ({inner_opt}): {bad} (all: {errs})"
    return errs


def check_field_eval(K, W, L, out, n, T, seed=0):
    theta = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    theta = torch.stack([theta * (1 + 0.05 * t) for t in range(T)])
    coords = torch.rand(T, n, 2, generator=torch.Generator().manual_seed(seed + 3)) * 2 - 1
    net = K.net_desc(L, W, out, OMEGA, True, nonlinearity=ACT)
    u = K.field_eval(net, _dev(K, theta), _dev(K, coords), T)
    from oracle.field import nf_apply
    for t in range(T):
        params = unflatten_params(theta[t].double(), [W] * L, 2, out, True)
        want = nf_apply(params, out, coords[t].double(), OMEGA, ACT)
        e = rel_l2(u[t].cpu(), want)
        assert e < TOL, f"field_eval W={W} out={out}: {e}"


# ------------------------------------------------------------------------------------------------------------
# Outer train steps: the trainers of meta_pde_b200/train.py (rollout kernels +
# the fused clip/Adam kernel mpde_outer_step) against oracle/meta.py's restatement of
#   src/maml_pde.py:203-226 (quirk 1: the norm of the MODEL grads rescales the lr-tree grads too; flax Adam on theta,
#   jax.experimental.optimizers.adam on alpha), src/leap_pde.py:175-195, src/nn_pde_maml.py:140-171.
# Compared after every step: theta, the alpha tree, Adam moments, meta_grad_norm.  Tolerance 1e-4 (rel-L2); parameter
# UPDATES (theta_s - theta_0 etc.) with the float32-resolution rule of _tol.
# ------------------------------------------------------------------------------------------------------------
def _upd_tol(x_new, x_old):
    d = (x_new.double() - x_old.double()).norm().item()
    return max(TOL, 8 * 6e-8 * x_new.double().norm().item() / max(d, 1e-300))


def _check_state(name, got, want, got0, errs, bad):
    """absolute agreement of the state + agreement of its update since the start"""
    e = rel_l2(got.cpu(), want)
    errs[name] = max(errs.get(name, 0.0), e)
    if not e < TOL:
        bad[name] = e
    if got0 is not None:
        eu = rel_l2(got.cpu().double() - got0.double(), want - got0.double())
        errs[name + "_update"] = max(errs.get(name + "_update", 0.0), eu)
        if not eu < _upd_tol(want, got0):
            bad[name + "_update"] = eu


def check_maml_train_steps(K, pde, W, L, n, T, K_steps=2, steps=3, inner_lr=1e-2, outer_lr=1e-3, lr_inner_lr=0.5,
                           grad_clip=1e14, inner_clip=1e14, optimizer="adam", seed=0, use_cuda_graph=False, resync=False):
    """resync=False: trainer and oracle run free for `steps` steps from the same start.  resync=True: after every step the
    oracle's state (theta, alpha tree, Adam moments) is set to the trainer's, so each step is compared from IDENTICAL
    inputs.  Why both: the SIREN meta-gradient is ill-conditioned in theta (a 1.3e-6 relative difference in theta moves
    ||g|| by 1.8e-4 at 3x64, omega 30), so a free-running float32 trajectory leaves the float64 one at ~10x per step
    whatever the kernels do: measured on the emulator at 3x64 / 128 points / clip 1e-3, step 1 agrees to 4.6e-7, step 3 of
    the free run to 1.8e-4 with glibc's 0.5-ulp sinf/cosf and to 1.8e-4 with compat.cuh's sincos alike."""
    from meta_pde_b200 import train
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    P = theta0.numel()
    tr = train.MamlTrainer(pde, L, W, theta0, bsize=T, inner_steps=K_steps, inner_lr=inner_lr, outer_lr=outer_lr,
                           lr_inner_lr=lr_inner_lr, inner_grad_clip=inner_clip, grad_clip=grad_clip, optimizer=optimizer,
                           kernels=K, use_cuda_graph=use_cuda_graph)
    lf = oracle_loss_fn(pde, W, L, 1.0)
    ocfg = ometa.MamlCfg(K_steps, inner_lr, inner_clip, True, 0.1)
    b1, b2 = {"adam": (0.9, 0.99), "rmsprop": (0.0, 0.8), "sgd": (0.9, 0.99)}[optimizer]
    opt_theta = ometa.FlaxAdam(outer_lr, b1, b2) if optimizer != "sgd" else None
    opt_lrs = ometa.JaxAdam(lr_inner_lr, 0.9, 0.99)
    o_theta, o_lrs = theta0.double(), torch.ones(K_steps, P, dtype=torch.float64)
    st_theta = opt_theta.init(o_theta) if opt_theta else None
    st_lrs = opt_lrs.init(o_lrs)
    errs, bad, clipped = {}, {}, []
    lrs0 = torch.ones(K_steps, P)
    for step in range(steps):
        batch = train.synthetic_maml_batch(pde, T, K_steps, n, n, seed=seed + 11 + step)
        inner, outer, final, aux = batch.views()
        tasks = [ometa.MamlTask(oracle_params(pde, batch.pde_params, t), [oracle_points(pde, s, t) for s in inner],
                                [oracle_points(pde, s, t) for s in outer], oracle_points(pde, final, t),
                                oracle_points(pde, aux, t)) for t in range(T)]
        losses, meta, auxt, gnorm = tr.train_step(step, batch.to(K.device))
        if opt_theta:
            o_theta, o_lrs, st_theta, st_lrs, o_losses, o_meta, o_gn = ometa.maml_train_step(
                step, ocfg, lf, o_theta, o_lrs, tasks, opt_theta, st_theta, opt_lrs, st_lrs, grad_clip)
        else:   # "sgd": Momentum(beta 0) on theta (trainer_util.py:606-608)
            (g_th, g_l), o_losses, o_meta = ometa.maml_multi_task_grad_and_losses(ocfg, lf, o_theta, tasks, o_lrs)
            o_gn = ometa.global_norm(g_th)
            g_th, g_l = ometa.clip_by_norm([g_th, g_l], o_gn, grad_clip)
            o_theta = o_theta - outer_lr * g_th
            o_lrs, st_lrs = opt_lrs.apply_at(step, o_lrs, g_l, st_lrs)
        o_theta, o_lrs = o_theta.detach(), o_lrs.detach()
        clipped.append(bool(o_gn > grad_clip))
        _check_state("theta", tr.theta, o_theta, theta0, errs, bad)
        _check_state("inner_lrs", tr.inner_lrs, o_lrs, lrs0, errs, bad)
        _check_state("meta_grad_norm", gnorm, o_gn.detach().reshape(1), None, errs, bad)
        _check_state("losses", losses, o_losses, None, errs, bad)
        _check_state("meta_losses", meta, o_meta[0], None, errs, bad)
        if opt_theta:
            _check_state("m_theta", tr.m_theta, st_theta["m"].detach(), None, errs, bad)
            _check_state("v_theta", tr.v_theta, st_theta["v"].detach(), None, errs, bad)
        _check_state("m_lrs", tr.m_lrs, st_lrs["m"].detach(), None, errs, bad)
        _check_state("v_lrs", tr.v_lrs, st_lrs["v"].detach(), None, errs, bad)
        if resync:
            f64 = lambda x, like: x.detach().cpu().double().reshape(like.shape)
            o_theta, o_lrs = f64(tr.theta, o_theta), f64(tr.inner_lrs, o_lrs)
            if opt_theta:
                st_theta = dict(st_theta, m=f64(tr.m_theta, st_theta["m"]), v=f64(tr.v_theta, st_theta["v"]))
            st_lrs = dict(st_lrs, m=f64(tr.m_lrs, st_lrs["m"]), v=f64(tr.v_lrs, st_lrs["v"]))
    assert not bad, f"MAML train_step {pde} {L}x{W} ({optimizer}, clip {grad_clip}): {bad} (all: {errs})"
    return errs, clipped


def check_leap_train_steps(K, pde, W, L, n, T, K_steps=3, steps=3, inner_lr=1e-3, outer_lr=1e-3, grad_clip=1e14, seed=0):
    from meta_pde_b200 import train
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    tr = train.LeapTrainer(pde, L, W, theta0, bsize=T, inner_steps=K_steps, inner_lr=inner_lr, outer_lr=outer_lr,
                           grad_clip=grad_clip, kernels=K)
    lf = oracle_loss_fn(pde, W, L, 1.0)
    ocfg = ometa.LeapCfg(K_steps, inner_lr, 1e14, True, True, True, "adam")
    opt = ometa.FlaxAdam(outer_lr, 0.9, 0.99)
    o_theta = theta0.double()
    st = opt.init(o_theta)
    errs, bad, clipped = {}, {}, []
    for step in range(steps):
        batch = train.synthetic_point_batch(pde, T, 1 + 2 * K_steps, n, seed=seed + 21 + step)
        g = batch.views()
        tasks = [ometa.LeapTask(oracle_params(pde, batch.pde_params, t), oracle_points(pde, g[0], t),
                                [oracle_points(pde, s, t) for s in g[1:1 + K_steps]],
                                [oracle_points(pde, s, t) for s in g[1 + K_steps:]]) for t in range(T)]
        losses, gnorm = tr.train_step(batch.to(K.device))
        o_theta, st, o_losses, o_gn = ometa.leap_train_step(ocfg, lf, o_theta, tasks, opt, st, grad_clip)
        o_theta = o_theta.detach()
        clipped.append(bool(o_gn > grad_clip))
        _check_state("theta", tr.theta, o_theta, theta0, errs, bad)
        _check_state("meta_grad_norm", gnorm, o_gn.detach().reshape(1), None, errs, bad)
        _check_state("losses", losses, o_losses, None, errs, bad)
        _check_state("m_theta", tr.m_theta, st["m"].detach(), None, errs, bad)
        _check_state("v_theta", tr.v_theta, st["v"].detach(), None, errs, bad)
    assert not bad, f"LEAP train_step {pde} {L}x{W} (clip {grad_clip}): {bad} (all: {errs})"
    return errs, clipped


def check_finetune_train_steps(K, pde, W, L, n, T, steps=3, outer_lr=1e-3, grad_clip=1e14, seed=0):
    from meta_pde_b200 import train
    out = PDE_OUT_DIM[pde]
    theta0 = _theta([W] * L, out, OMEGA, OMEGA, True, seed=seed, bias_scale=0.05)
    tr = train.FinetuneTrainer(pde, L, W, theta0, bsize=T, outer_lr=outer_lr, grad_clip=grad_clip, kernels=K)
    lf = oracle_loss_fn(pde, W, L, 1.0)
    opt = ometa.FlaxAdam(outer_lr, 0.9, 0.99)
    o_theta = theta0.double()
    st = opt.init(o_theta)
    errs, bad, clipped = {}, {}, []
    for step in range(steps):
        batch = train.synthetic_point_batch(pde, T, 1, n, seed=seed + 31 + step)
        (sets,) = batch.views()
        ob = [(oracle_points(pde, sets, t), oracle_params(pde, batch.pde_params, t)) for t in range(T)]
        loss, terms, gnorm = tr.train_step(batch.to(K.device))
        o_theta, st, o_loss, o_aux, o_gn = ometa.finetune_train_step(lf, o_theta, ob, opt, st, grad_clip)
        o_theta = o_theta.detach()
        clipped.append(bool(o_gn > grad_clip))
        _check_state("theta", tr.theta, o_theta, theta0, errs, bad)
        _check_state("grad_norm", gnorm, o_gn.detach().reshape(1), None, errs, bad)
        _check_state("loss", loss.sum().reshape(1), o_loss.reshape(1), None, errs, bad)
        _check_state("loss_aux", terms.mean(dim=0), torch.stack([o_aux[k] for k in PDE_TERMS[pde]]), None, errs, bad)
        _check_state("m_theta", tr.m_theta, st["m"].detach(), None, errs, bad)
        _check_state("v_theta", tr.v_theta, st["v"].detach(), None, errs, bad)
    assert not bad, f"fine-tune train_step {pde} {L}x{W} (clip {grad_clip}): {bad} (all: {errs})"
    return errs, clipped
