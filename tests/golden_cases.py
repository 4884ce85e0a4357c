"""Checks against tests/golden/ref_*.npz -- outputs of the REFERENCE'S OWN SOURCE executed under the jax/flax
stand-in (tests/golden/make_reference_fixtures.py).  Two consumers: the oracle (float64, must agree to ~1e-9:
this is what pins the oracle) and the kernels through the C ABI (float32, north-star tolerance 1e-4)."""
import glob
import os

import numpy as np
import torch

from helpers import oracle_params, oracle_points, rel_l2
from meta_pde_b200._abi import PDE_OUT_DIM, PDE_SETS, PDE_TERMS
from oracle import meta as ometa
from oracle import pdes as opdes
from oracle.field import make_field_fn, unflatten_params

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_KERNEL = 1e-4    # BASELINE.json north_star
TOL_ORACLE = 1e-9    # float64 restatement vs float64 run of the reference source


def fixtures(kind):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, f"ref_{kind}_*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    d = {k: torch.from_numpy(np.asarray(z[k])) for k in z.files}
    parts = name.split("_")
    kind = parts[1]
    pde = "_".join(parts[2:]).split("_2x")[0].split("_3x")[0]
    d["_pde"], d["_kind"] = pde, kind
    d["_act"] = "swish" if int(d.get("act", torch.tensor(0))) else "sin"   # fixtures of the `--nosiren` network carry act = 1
    return d


def _sets(d, prefix, pde):
    return [d[f"{prefix}set{s}"] for s in range(len(PDE_SETS[pde]))]


def _assert(errs, tol, what):
    bad = {k: e for k, e in errs.items() if not (e < tol)}
    assert not bad, f"{what}: rel errors above {tol}: {bad} (all: {errs})"


# ------------------------------------------------------------------------------------------- oracle side
def oracle_vs_loss_fixture(name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out = (int(x) for x in d["meta"])
    lf = ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, float(d["omega"]), True, out, d["_act"]), float(d["bc_weight"]))
    sets = _sets(d, "", pde)
    errs = {}
    for t in range(T):
        th = d["theta"].double().clone().requires_grad_(True)
        pts, prm = oracle_points(pde, sets, t), oracle_params(pde, d["pde_params"], t)
        l, aux = lf(th, pts, prm)
        (g,) = torch.autograd.grad(l, th)
        ffn = make_field_fn(unflatten_params(d["theta"].double(), [W] * L, 2, out, True), out, float(d["omega"]), d["_act"])
        if pde == "poisson":
            q = opdes.poisson_residual(ffn, pts[1], prm[0])
        elif pde == "td_burgers":
            q = opdes.burgers_loss_domain_fn(ffn, pts[3], prm)
        else:
            q = opdes.hyper_energy_density(ffn, pts[5], prm)
        for k, got, want in [("loss", l, d["loss"][t]), ("grad", g, d["grad"][t]), ("domain_quantity", q, d["domain_quantity"][t]),
                             ("terms", torch.stack([aux[n] for n in PDE_TERMS[pde]]), d["terms"][t])]:
            errs[k] = max(errs.get(k, 0.0), rel_l2(got.detach(), want))
    _assert(errs, TOL_ORACLE, f"oracle vs {name}")


def _maml_tasks(d, pde, T, K):
    tasks = []
    for t in range(T):
        tasks.append(ometa.MamlTask(oracle_params(pde, d["pde_params"], t),
                                    [oracle_points(pde, _sets(d, f"inner{k}_", pde), t) for k in range(K)],
                                    [oracle_points(pde, _sets(d, f"outer{k}_", pde), t) for k in range(K)],
                                    oracle_points(pde, _sets(d, "final_", pde), t), oracle_points(pde, _sets(d, "aux_", pde), t)))
    return tasks


def oracle_vs_maml_fixture(name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out, K = (int(x) for x in d["meta"])
    inner_lr, clip, decay = (float(x) for x in d["hyper"])
    lf = ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, float(d["omega"]), True, out, d["_act"]), float(d["bc_weight"]))
    cfg = ometa.MamlCfg(K, inner_lr, clip, True, decay)
    (g_theta, g_lrs), losses, (meta, aux) = ometa.maml_multi_task_grad_and_losses(
        cfg, lf, d["theta"].double(), _maml_tasks(d, pde, T, K), d["inner_lrs"].double())
    errs = {"g_theta": rel_l2(g_theta, d["g_theta"]), "g_lrs": rel_l2(g_lrs, d["g_lrs"]), "losses": rel_l2(losses, d["losses"]),
            "meta_losses": rel_l2(meta, d["meta_losses"]),
            "aux_terms": rel_l2(torch.stack([aux[n] for n in PDE_TERMS[pde]], dim=1), d["aux_terms"])}
    _assert(errs, TOL_ORACLE, f"oracle vs {name}")


def _leap_tasks(d, pde, T, K):
    return [ometa.LeapTask(oracle_params(pde, d["pde_params"], t), oracle_points(pde, _sets(d, "loss0_", pde), t),
                           [oracle_points(pde, _sets(d, f"step{k}_", pde), t) for k in range(K)],
                           [oracle_points(pde, _sets(d, f"new{k}_", pde), t) for k in range(K)]) for t in range(T)]


def oracle_vs_leap_fixture(name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out, K = (int(x) for x in d["meta"])
    inner_lr, clip = (float(x) for x in d["hyper"])
    lf = ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, float(d["omega"]), True, out, d["_act"]), float(d["bc_weight"]))
    cfg = ometa.LeapCfg(K, inner_lr, clip, True, True, True, "adam")
    mg, losses = ometa.leap_multi_task_grad_and_losses(cfg, lf, d["theta"].double(), _leap_tasks(d, pde, T, K))
    errs = {"meta_grad": rel_l2(mg, d["meta_grad"]), "losses": rel_l2(losses, d["losses"])}
    # Adam's m / (sqrt(v) + eps) at step 1 amplifies last-bit differences of tiny gradient entries: 1e-7 here
    _assert(errs, 1e-7, f"oracle vs {name}")


# ------------------------------------------------------------------------------------------- kernel side
def _dev(K, x):
    return x.to(torch.float32).contiguous().to(K.device)


def kernels_vs_loss_fixture(K, name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out = (int(x) for x in d["meta"])
    net = K.net_desc(L, W, out, float(d["omega"]), True, nonlinearity=d["_act"])
    sets = [_dev(K, s) for s in _sets(d, "", pde)]
    bw = float(d["bc_weight"])
    loss, terms, res = K.loss(net, pde, bw, _dev(K, d["theta"]), sets, _dev(K, d["pde_params"]), T, want_residual=True)
    loss1, terms1, grad = K.loss_grad(net, pde, bw, _dev(K, d["theta"]), sets, _dev(K, d["pde_params"]), T)
    q = res.cpu().double()
    if pde == "td_burgers":
        q = q ** 2           # the reference's loss_domain_fn returns the squared residual
    errs = {"loss": rel_l2(loss.cpu(), d["loss"]), "loss1": rel_l2(loss1.cpu(), d["loss"]), "terms": rel_l2(terms.cpu(), d["terms"]),
            "grad": rel_l2(grad.cpu(), d["grad"]), "domain_quantity": rel_l2(q, d["domain_quantity"])}
    _assert(errs, TOL_KERNEL, f"kernels vs {name}")


def kernels_vs_maml_fixture(K, name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out, Ks = (int(x) for x in d["meta"])
    inner_lr, clip, decay = (float(x) for x in d["hyper"])
    net = K.net_desc(L, W, out, float(d["omega"]), True, nonlinearity=d["_act"])
    cfg = dict(inner_steps=Ks, inner_lr=inner_lr, inner_grad_clip=clip, softplus_lrs=True, outer_loss_decay=decay)
    S = lambda pre: [_dev(K, s) for s in _sets(d, pre, pde)]
    got = K.maml_outer(net, pde, float(d["bc_weight"]), cfg, _dev(K, d["theta"]), _dev(K, d["inner_lrs"]),
                       [S(f"inner{k}_") for k in range(Ks)], [S(f"outer{k}_") for k in range(Ks)], S("final_"), S("aux_"),
                       _dev(K, d["pde_params"]), T)
    errs = {k: rel_l2(got[k].cpu(), d[k]) for k in ("g_theta", "g_lrs", "losses", "meta_losses", "aux_terms")}
    _assert(errs, TOL_KERNEL, f"kernels vs {name}")


def kernels_vs_leap_fixture(K, name):
    d = load(name)
    pde = d["_pde"]
    W, L, T, out, Ks = (int(x) for x in d["meta"])
    inner_lr, clip = (float(x) for x in d["hyper"])
    net = K.net_desc(L, W, out, float(d["omega"]), True, nonlinearity=d["_act"])
    cfg = dict(inner_steps=Ks, inner_lr=inner_lr, inner_grad_clip=clip, norm=True, loss_in_distance=True, stabilize=True,
               inner_opt="adam", beta1=0.9, beta2=0.99, eps=1e-8)
    S = lambda pre: [_dev(K, s) for s in _sets(d, pre, pde)]
    got = K.leap_outer(net, pde, float(d["bc_weight"]), cfg, _dev(K, d["theta"]), S("loss0_"),
                       [S(f"step{k}_") for k in range(Ks)], [S(f"new{k}_") for k in range(Ks)], _dev(K, d["pde_params"]), T)
    errs = {"meta_grad": rel_l2(got["meta_grad"].cpu(), d["meta_grad"]), "losses": rel_l2(got["losses"].cpu(), d["losses"])}
    _assert(errs, TOL_KERNEL, f"kernels vs {name}")
