#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN SOURCE FILES (read from /root/reference,
never copied) under the torch-backed jax/flax stand-in of tests/golden/jaxshim.py.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_reference_fixtures.py

What runs from the reference, unmodified:
    src/nets/field.py        NeuralPotential / make_nf_ndim(2) .partial(...).init_by_shape(...), nf_apply, laplace_operator
    src/poisson/poisson_common.py            loss_fn (+ boundary_conditions, source)
    src/burgers/td_burgers_common.py         loss_fn, loss_domain_fn (+ burgers_formulation/default.py)
    src/elasticity/hyper_elasticity_common.py loss_fn, loss_domain_fn, deformation_gradient, right_cauchygreen
    src/nets/maml.py         multi_task_grad_and_losses -> single_task_grad_and_losses -> single_task_rollout -> maml_inner_step
    src/nets/leap.py         multi_task_grad_and_losses -> ... -> leap_inner_step, get_meta_grad_increment
What is restated here (driver glue that lives inside main() of the reference's absl apps and cannot be imported):
    the loss wrapper and task closures of src/maml_pde.py:76-98 / src/leap_pde.py:66-85, MamlDef / LeapDef
    construction (src/maml_pde.py:100-110, src/leap_pde.py:87-97), with the point SAMPLERS replaced by a
    recorder-backed synthetic sampler (sampler outputs are inputs of the hot path).

Every fixture stores the arrays the reference code consumed (theta in the C-ABI flat layout, the point sets in
ABI order, pde_params) and the arrays it produced (loss, terms, grad; MAML meta-grads / losses / meta-losses /
aux; LEAP meta-grad / losses), in float64.
"""
import math
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import numpy as onp  # noqa: E402
import torch  # noqa: E402

import jaxshim  # noqa: E402

jax, flax = jaxshim.install()
sys.path.insert(0, REF)

from absl import flags  # noqa: E402
from src.util import common_flags  # noqa: E402,F401  (the reference's flag definitions)

# driver-local flags read inside src/nets/maml.py / leap.py (defined in src/maml_pde.py:50-58)
flags.DEFINE_float("inner_grad_clip", 1e14, "inner grad clipping")
FLAGS = flags.FLAGS
FLAGS(["fixtures", "--siren_omega=30.0", "--siren_omega0=30.0", "--bc_weight=1.0", "--burgers_pde=default"])

import jax.numpy as np  # noqa: E402  (the shim)
from src.nets import field as ref_field  # noqa: E402
from src.nets import maml as ref_maml  # noqa: E402
from src.nets import leap as ref_leap  # noqa: E402
from src.poisson import poisson_common as ref_poisson  # noqa: E402
from src.burgers import td_burgers_common as ref_burgers  # noqa: E402
from src.elasticity import hyper_elasticity_common as ref_hyper  # noqa: E402

from meta_pde_b200._abi import PDE_NPRM, PDE_OUT_DIM, PDE_SETS, PDE_TERMS  # noqa: E402  (names / orders only)

PDE_MOD = {"poisson": ref_poisson, "td_burgers": ref_burgers, "hyper_elasticity": ref_hyper}
BASE_FIELD = {"poisson": ref_field.NeuralPotential, "td_burgers": ref_field.NeuralPotential,
              "hyper_elasticity": ref_field.make_nf_ndim(2)}   # *_def.py: BaseField
DT = torch.float64


# ---------------------------------------------------------------------------------- theta <-> flax params
def flatten_model_tree(tree, L):
    """Flax param dict (or a gradient / lr tree with the same keys) -> ABI flat order; leading dims kept."""
    parts = []
    for l in range(L + 1):
        d = tree[f"Dense_{l}"]
        k, b = d["kernel"], d["bias"]
        lead = k.shape[:-2]
        parts += [k.reshape(*lead, -1), b.reshape(*lead, -1)]
    for name in ("log_input_scale", "log_output_scale"):
        if name in tree:
            x = tree[name]
            parts.append(x.reshape(*x.shape[:-2], -1))
    return torch.cat(parts, dim=-1)


def unflatten_like(flat, params, L):
    """ABI flat [.., P] -> a dict shaped like `params` (leading dims kept)."""
    out, off = {}, 0
    lead = flat.shape[:-1]

    def take(shape):
        nonlocal off
        n = int(onp.prod(shape))
        x = flat[..., off:off + n].reshape(*lead, *shape)
        off += n
        return x

    for l in range(L + 1):
        d = params[f"Dense_{l}"]
        k = take(tuple(d["kernel"].shape))
        b = take(tuple(d["bias"].shape))
        out[f"Dense_{l}"] = {"kernel": k, "bias": b}
    for name in ("log_input_scale", "log_output_scale"):
        if name in params:
            out[name] = take(tuple(params[name].shape))
    assert off == flat.shape[-1]
    return out


# ---------------------------------------------------------------------------------- synthetic samplers
def sample_params(pde, key):
    g = jaxshim._gen(key, 0x61)
    if pde == "poisson":
        return (torch.as_tensor(g.standard_normal((2, 3)), dtype=DT), torch.as_tensor(g.uniform(-1, 1, 5), dtype=DT),
                torch.zeros(2, dtype=DT))
    if pde == "td_burgers":
        return (torch.as_tensor(g.uniform(80, 100, 1), dtype=DT), torch.as_tensor(g.uniform(-2, 2, 2), dtype=DT))
    return (torch.as_tensor(g.uniform(0.25, 0.75, 2), dtype=DT), torch.as_tensor(g.uniform(0.9, 1.1, 2), dtype=DT),
            torch.zeros(25, 5, dtype=DT), 25)


def params_to_abi(pde, p):
    if pde == "poisson":
        return torch.cat([p[0].reshape(-1), p[1], p[2]])
    if pde == "td_burgers":
        return torch.cat([p[0], p[1]])
    return torch.cat([p[0], p[1]])


def sample_points(pde, key, n):
    """-> the tuple pde.loss_fn takes.  Geometry as in the reference's samplers (circle boundary / unit disc,
    (x,t) in [0,1]^2, unit square), set sizes free."""
    g = jaxshim._gen(key, 0x62)
    u = lambda *s: torch.as_tensor(g.random(s), dtype=DT)
    if pde == "poisson":
        th = 2 * math.pi * u(n)
        bdy = torch.stack([torch.cos(th), torch.sin(th)], 1)
        r, ph = torch.sqrt(u(n)), 2 * math.pi * u(n)
        dom = torch.stack([r * torch.cos(ph), r * torch.sin(ph)], 1)
        return (bdy, dom)
    if pde == "td_burgers":
        nb = max(n // 4, 3)
        tl = u(nb)
        left = torch.stack([torch.zeros(nb, dtype=DT), tl], 1)
        right = torch.stack([torch.ones(nb, dtype=DT), tl], 1)
        init = torch.stack([u(n + 2), torch.zeros(n + 2, dtype=DT)], 1)
        return (left, right, init, u(n, 2))
    top = torch.stack([u(n), torch.ones(n, dtype=DT)], 1)
    bot = torch.stack([u(n), torch.zeros(n, dtype=DT)], 1)
    side = torch.stack([torch.zeros(3, dtype=DT), u(3)], 1)
    return (top, bot, side, side, side, u(n, 2))


def points_to_abi(pde, pts):
    if pde == "hyper_elasticity":
        return [pts[0], pts[1], pts[5]]
    return list(pts)


# ---------------------------------------------------------------------------------- driver glue (restated)
def make_driver(pde):
    mod = PDE_MOD[pde]

    def loss_fn(field_fn, points, params):           # src/maml_pde.py:76-85
        boundary_losses, domain_losses = mod.loss_fn(field_fn, points, params)
        loss = np.sum(FLAGS.bc_weight * np.sum(np.array([bl for bl in boundary_losses.values()])) +
                      np.sum(np.array([dl for dl in domain_losses.values()])))
        return loss, {**boundary_losses, **domain_losses}

    return loss_fn


ACT = "sin"   # "swish": the `--nosiren` network (nonlinearity=nn.swish, src/maml_pde.py:115); set by main()


def _sfx():
    return "" if ACT == "sin" else "_" + ACT


def build_model(pde, W, L, seed, bias_scale=0.05):
    Field = BASE_FIELD[pde].partial(sizes=[W] * L, dense_args=(), nonlinearity=np.sin if ACT == "sin" else flax.nn.swish,
                                    omega=FLAGS.siren_omega,
                                    omega0=FLAGS.siren_omega0, log_scale=FLAGS.log_scale)      # src/maml_pde.py:112-119
    key, subkey = jax.random.split(jax.random.PRNGKey(seed))
    _, init_params = Field.init_by_shape(subkey, [((1, 2), np.float32)])                        # src/maml_pde.py:121-124
    g = onp.random.default_rng(seed + 17)
    for l in range(L + 1):   # the reference's zero biases would leave the bias paths untested
        b = init_params[f"Dense_{l}"]["bias"]
        init_params[f"Dense_{l}"]["bias"] = b + bias_scale * torch.as_tensor(g.standard_normal(tuple(b.shape)), dtype=DT)
    return Field, flax.nn.Model(Field, init_params)


class Recorder:
    def __init__(self):
        self.calls = []


def save(name, **arrays):
    out = {} if ACT == "sin" else {"act": onp.array(1)}   # SIREN fixtures keep their round-1/2 key set
    for k, v in arrays.items():
        if torch.is_tensor(v):
            v = v.detach().cpu().numpy()
        out[k] = onp.asarray(v)
    path = os.path.join(HERE, name + ".npz")
    onp.savez_compressed(path, **out)
    print("wrote", path, {k: tuple(v.shape) for k, v in out.items()})


def stack_sets(list_of_abi_sets):
    """[task][set] tensors -> [set] of [T, n, 2]"""
    n_sets = len(list_of_abi_sets[0])
    return [torch.stack([t[s] for t in list_of_abi_sets]) for s in range(n_sets)]


# ---------------------------------------------------------------------------------- fixture kinds
def fixture_loss_grad(pde, W, L, n, T, seed):
    """pde.loss_fn + driver wrapper + jax.value_and_grad wrt the Model (as maml_inner_step does, maml.py:66-70)."""
    loss_fn = make_driver(pde)
    Field, model = build_model(pde, W, L, seed)
    theta = flatten_model_tree(model.params, L)
    keys = jax.random.split(jax.random.PRNGKey(seed + 1), T)
    losses, terms, grads, sets, prms, resid = [], [], [], [], [], []
    for t in range(T):
        k1, k2 = jax.random.split(keys[t])
        params = sample_params(pde, k1)
        pts = sample_points(pde, k2, n)
        (loss, aux), grad = jax.value_and_grad(lambda key, m: loss_fn(m, pts, params), argnums=1, has_aux=True)(None, model)
        losses.append(loss.detach())
        terms.append(torch.stack([aux[k].detach() for k in PDE_TERMS[pde]]))
        grads.append(flatten_model_tree(grad.params, L).detach())
        sets.append(points_to_abi(pde, pts))
        prms.append(params_to_abi(pde, params))
        # per-point domain quantity straight from the reference functions
        if pde == "poisson":
            r = ref_field.vmap_laplace_operator(pts[1], model, lambda x: 1 + 0.1 * model(x) ** 2) - \
                ref_poisson.vmap_source(pts[1], params[0])
        elif pde == "td_burgers":
            r = ref_burgers.loss_domain_fn(model, pts[3], params)       # squared residual
        else:
            r = ref_hyper.loss_domain_fn(model, pts[5], params)         # energy density
        resid.append(r.detach())
    arrays = dict(theta=theta.detach(), pde_params=torch.stack(prms), loss=torch.stack(losses), terms=torch.stack(terms),
                  grad=torch.stack(grads), domain_quantity=torch.stack(resid),
                  meta=onp.array([W, L, T, PDE_OUT_DIM[pde]]), omega=FLAGS.siren_omega, bc_weight=FLAGS.bc_weight)
    for s, x in enumerate(stack_sets(sets)):
        arrays[f"set{s}"] = x
    save(f"ref_loss_{pde}_{L}x{W}{_sfx()}", **arrays)


def fixture_maml(pde, W, L, n, T, K, seed, inner_lr, clip, decay):
    loss_fn = make_driver(pde)
    Field, model = build_model(pde, W, L, seed)
    theta = flatten_model_tree(model.params, L)
    P = theta.numel()
    rec = Recorder()
    FLAGS.inner_grad_clip = clip

    def make_task_loss_fns(key):                     # src/maml_pde.py:87-98 with a recording sampler
        params = sample_params(pde, key)
        rec.calls.append(("params", params))

        def inner_loss(key, field_fn, params=params):
            pts = sample_points(pde, key, n)
            rec.calls.append(("inner", pts))
            return loss_fn(field_fn, pts, params)

        def outer_loss(key, field_fn, params=params):
            pts = sample_points(pde, key, n)
            rec.calls.append(("outer", pts))
            return loss_fn(field_fn, pts, params)

        return inner_loss, outer_loss

    maml_def = ref_maml.MamlDef(make_inner_opt=flax.optim.Momentum(learning_rate=inner_lr, beta=0.0).create,
                                make_task_loss_fns=make_task_loss_fns, inner_steps=K, n_batch_tasks=T,
                                softplus_lrs=True, outer_loss_decay=decay)          # src/maml_pde.py:100-110
    g = onp.random.default_rng(seed + 5)
    lrs_flat = torch.as_tensor(1.0 + 0.3 * g.standard_normal((K, P)), dtype=DT)
    inner_lrs = flax.nn.Model(Field, unflatten_like(lrs_flat, model.params, L))     # [K, *leaf.shape] per leaf (maml_pde.py:151-154)
    grads, losses, meta_losses = ref_maml.multi_task_grad_and_losses(maml_def, jax.random.PRNGKey(seed + 2), model, inner_lrs)
    g_model, g_lrs = grads
    meta, aux = meta_losses
    # replay the recorder: per task [params, (inner, outer) x K, inner (final), outer (aux)]
    per_task = 2 * K + 3
    assert len(rec.calls) == T * per_task, (len(rec.calls), T, per_task)
    inner = [[] for _ in range(K)]
    outer = [[] for _ in range(K)]
    final, auxp, prms = [], [], []
    for t in range(T):
        c = rec.calls[t * per_task:(t + 1) * per_task]
        assert c[0][0] == "params"
        prms.append(params_to_abi(pde, c[0][1]))
        for k in range(K):
            assert c[1 + 2 * k][0] == "inner" and c[2 + 2 * k][0] == "outer"
            inner[k].append(points_to_abi(pde, c[1 + 2 * k][1]))
            outer[k].append(points_to_abi(pde, c[2 + 2 * k][1]))
        assert c[2 * K + 1][0] == "inner" and c[2 * K + 2][0] == "outer"
        final.append(points_to_abi(pde, c[2 * K + 1][1]))
        auxp.append(points_to_abi(pde, c[2 * K + 2][1]))
    arrays = dict(theta=theta.detach(), inner_lrs=lrs_flat, pde_params=torch.stack(prms),
                  g_theta=flatten_model_tree(g_model.params, L).detach(), g_lrs=flatten_model_tree(g_lrs.params, L).detach(),
                  losses=losses.detach(), meta_losses=meta.detach(),
                  aux_terms=torch.stack([aux[k].detach() for k in PDE_TERMS[pde]], dim=1),
                  meta=onp.array([W, L, T, PDE_OUT_DIM[pde], K]), omega=FLAGS.siren_omega, bc_weight=FLAGS.bc_weight,
                  hyper=onp.array([inner_lr, clip, decay]))
    for k in range(K):
        for s, x in enumerate(stack_sets(inner[k])):
            arrays[f"inner{k}_set{s}"] = x
        for s, x in enumerate(stack_sets(outer[k])):
            arrays[f"outer{k}_set{s}"] = x
    for s, x in enumerate(stack_sets(final)):
        arrays[f"final_set{s}"] = x
    for s, x in enumerate(stack_sets(auxp)):
        arrays[f"aux_set{s}"] = x
    tag = "clip" if clip < 1e10 else "noclip"
    save(f"ref_maml_{pde}_{L}x{W}_{tag}{_sfx()}", **arrays)
    FLAGS.inner_grad_clip = 1e14


def fixture_leap(pde, W, L, n, T, K, seed, inner_lr, clip):
    loss_fn = make_driver(pde)
    Field, model = build_model(pde, W, L, seed)
    theta = flatten_model_tree(model.params, L)
    rec = Recorder()
    FLAGS.inner_grad_clip = clip

    def make_task_loss_fn(key):                      # src/leap_pde.py:75-85
        params = sample_params(pde, key)
        rec.calls.append(("params", params))

        def variational_energy_estimator(inner_key, field_fn, params=params):
            pts = sample_points(pde, inner_key, n)
            rec.calls.append(("pts", pts))
            return loss_fn(field_fn, pts, params)

        return variational_energy_estimator

    leap_def = ref_leap.LeapDef(make_inner_opt=flax.optim.Adam(learning_rate=inner_lr, beta1=0.9, beta2=0.99).create,
                                make_task_loss_fn=make_task_loss_fn, inner_steps=K, n_batch_tasks=T, norm=True,
                                loss_in_distance=True, stabilize=True)              # src/leap_pde.py:87-97
    grads, losses = ref_leap.multi_task_grad_and_losses(leap_def, jax.random.PRNGKey(seed + 3), model)
    loss_arr = losses[0]        # (losses, aux) tuple: leap.py:143-145
    per_task = 2 * K + 2        # params, loss0, (step, new) x K
    assert len(rec.calls) == T * per_task, (len(rec.calls), T, per_task)
    loss0, step, new, prms = [], [[] for _ in range(K)], [[] for _ in range(K)], []
    for t in range(T):
        c = rec.calls[t * per_task:(t + 1) * per_task]
        prms.append(params_to_abi(pde, c[0][1]))
        loss0.append(points_to_abi(pde, c[1][1]))
        for k in range(K):
            step[k].append(points_to_abi(pde, c[2 + 2 * k][1]))
            new[k].append(points_to_abi(pde, c[3 + 2 * k][1]))
    arrays = dict(theta=theta.detach(), pde_params=torch.stack(prms), meta_grad=flatten_model_tree(grads.params, L).detach(),
                  losses=loss_arr.detach(), meta=onp.array([W, L, T, PDE_OUT_DIM[pde], K]), omega=FLAGS.siren_omega,
                  bc_weight=FLAGS.bc_weight, hyper=onp.array([inner_lr, clip]))
    for s, x in enumerate(stack_sets(loss0)):
        arrays[f"loss0_set{s}"] = x
    for k in range(K):
        for s, x in enumerate(stack_sets(step[k])):
            arrays[f"step{k}_set{s}"] = x
        for s, x in enumerate(stack_sets(new[k])):
            arrays[f"new{k}_set{s}"] = x
    save(f"ref_leap_{pde}_{L}x{W}{_sfx()}", **arrays)
    FLAGS.inner_grad_clip = 1e14


def main_swish():
    """Row f4: the same reference functions with nonlinearity = flax.nn.swish (no omega, field.py:187-188)."""
    global ACT
    ACT = "swish"
    torch.manual_seed(0)
    for pde in ("poisson", "td_burgers", "hyper_elasticity"):
        fixture_loss_grad(pde, 32, 2, 24, 2, seed=41)
    fixture_loss_grad("poisson", 64, 3, 40, 1, seed=42)
    fixture_maml("poisson", 32, 2, 16, 2, 2, seed=51, inner_lr=1e-2, clip=1e14, decay=0.1)
    fixture_maml("td_burgers", 32, 2, 16, 2, 2, seed=52, inner_lr=1e-2, clip=1e-2, decay=0.5)
    fixture_leap("hyper_elasticity", 32, 2, 16, 2, 2, seed=61, inner_lr=1e-4, clip=1e14)
    ACT = "sin"


def main():
    if "--only-swish" in sys.argv[1:]:
        return main_swish()
    torch.manual_seed(0)
    for pde in ("poisson", "td_burgers", "hyper_elasticity"):
        fixture_loss_grad(pde, 32, 2, 24, 2, seed=11)
    fixture_loss_grad("poisson", 64, 3, 40, 1, seed=12)
    lr = {"poisson": 1e-2, "td_burgers": 1e-2, "hyper_elasticity": 1e-4}
    for pde in ("poisson", "td_burgers", "hyper_elasticity"):
        fixture_maml(pde, 32, 2, 16, 2, 2, seed=21, inner_lr=lr[pde], clip=1e14, decay=0.1)
    fixture_maml("poisson", 32, 2, 16, 2, 3, seed=22, inner_lr=1e-2, clip=1e-2, decay=0.5)
    fixture_leap("poisson", 32, 2, 16, 2, 3, seed=31, inner_lr=1e-3, clip=1e14)
    fixture_leap("td_burgers", 32, 2, 16, 2, 2, seed=32, inner_lr=1e-3, clip=0.5)
    # width 128: the tensor-core (tcgen05) kernels + tc_wgrad_kernel, pinned to the reference source directly (round 2)
    fixture_loss_grad("poisson", 128, 2, 40, 1, seed=13)
    fixture_loss_grad("hyper_elasticity", 128, 2, 24, 1, seed=14)
    fixture_loss_grad("td_burgers", 128, 2, 70, 1, seed=15)
    fixture_maml("poisson", 128, 2, 16, 2, 2, seed=23, inner_lr=1e-2, clip=1e14, decay=0.1)
    fixture_maml("td_burgers", 128, 2, 70, 2, 2, seed=24, inner_lr=1e-2, clip=1e-2, decay=0.1)
    fixture_leap("poisson", 128, 2, 16, 2, 2, seed=33, inner_lr=1e-3, clip=1e14)
    main_swish()


if __name__ == "__main__":
    main()
