"""Per-PDE plugin modules (mirror of src/{poisson,burgers,elasticity}/*_def.py + *_common.py and src/get_pde.py).

Each ``PdeDef`` exposes the attributes the reference's drivers use: ``sample_params(key)``,
``sample_points(key, n, params)``, ``sample_points_in_domain(key, n, params)``,
``loss_fn(field_fn, points, params) -> (boundary_losses: dict, domain_losses: dict)``, ``BaseField``, ``dim``,
``solve_fenics`` / ``plot_solution`` (unavailable: FEniCS / matplotlib are out of the path and absent).

``loss_fn`` recognises a SIREN ``Model`` and evaluates ALL terms in one fused jet kernel (mpde_loss); it is
differentiable to second order with torch autograd through hand-written adjoint / Hessian-vector kernels
(mpde_loss_grad, mpde_loss_hvp) -- the counterpart of wrapping XLA custom calls in nested custom_vjp.
Samplers run on the host with compat.random (threefry2x32) following the reference's key derivations.
"""
import math

import numpy as np
import torch

from .. import ops
from .._abi import PDE_SETS, PDE_TERMS
from . import field as cfield
from . import random as jr
from .flags import FLAGS

_BOUNDARY_TERMS = {"poisson": ["boundary_loss"], "td_burgers": ["loss_initial", "loss_left", "loss_right"],
                   "hyper_elasticity": ["loss_bottom", "loss_top"]}
_DOMAIN_TERMS = {"poisson": ["domain_loss"], "td_burgers": ["loss_domain"], "hyper_elasticity": ["loss_domain"]}


# ------------------------------------------------------------------------------------------ autograd bridge
class _Call:
    """Everything but theta of one loss evaluation (T = 1)."""

    def __init__(self, K, net, pde, sets, prm):
        self.K, self.net, self.pde, self.sets, self.prm = K, net, pde, sets, prm


class _GradFn(torch.autograd.Function):
    """theta -> d loss / d theta (mpde_loss_grad); its VJP is the Hessian-vector product (mpde_loss_hvp)."""

    @staticmethod
    def forward(ctx, theta, call, bc_weight):
        _, _, grad = call.K.loss_grad(call.net, call.pde, bc_weight, theta.detach().contiguous(), call.sets, call.prm, 1)
        ctx.call, ctx.bc_weight = call, bc_weight
        ctx.save_for_backward(theta)
        return grad[0].clone()

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, v):
        (theta,) = ctx.saved_tensors
        c = ctx.call
        _, _, _, hv = c.K.loss_hvp(c.net, c.pde, ctx.bc_weight, theta.detach().contiguous(), v.contiguous(), c.sets, c.prm, 1,
                                   want_grad=False)
        return hv[0].clone(), None, None


class _TermsFn(torch.autograd.Function):
    """theta -> the loss terms (mpde_loss).  Backward supports the reference's weighting
    (one weight for all boundary terms, one for all domain terms: src/maml_pde.py:76-85)."""

    @staticmethod
    def forward(ctx, theta, call):
        _, terms = call.K.loss(call.net, call.pde, 1.0, theta.detach().contiguous(), call.sets, call.prm, 1)
        ctx.call = call
        ctx.save_for_backward(theta)
        return terms[0].clone()

    @staticmethod
    def backward(ctx, g_terms):
        (theta,) = ctx.saved_tensors
        pde = ctx.call.pde
        names = PDE_TERMS[pde]
        g = g_terms.detach().tolist()
        gb = {g[names.index(n)] for n in _BOUNDARY_TERMS[pde]}
        gd = {g[names.index(n)] for n in _DOMAIN_TERMS[pde]}
        if len(gb) != 1 or len(gd) != 1 or next(iter(gd)) == 0.0:
            raise NotImplementedError("metapde_b200: loss terms must be combined as w_b * sum(boundary) + w_d * sum(domain), w_d != 0")
        wb, wd = next(iter(gb)), next(iter(gd))
        return wd * _GradFn.apply(theta, ctx.call, wb / wd), None


# ------------------------------------------------------------------------------------------ plugin objects
def _dev():
    # samplers are host-side plumbing: without CUDA their outputs stay on the host (compute entry points still raise)
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def _to_dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).to(_dev())


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def _unavailable(name):
    def f(*a, **k):
        raise RuntimeError(f"metapde_b200: {name} needs FEniCS / matplotlib, which are outside the hot path and not installed")
    return f


class PdeDef:
    name = None
    dim = 1
    BaseField = None
    solve_fenics = staticmethod(_unavailable("solve_fenics"))
    plot_solution = staticmethod(_unavailable("plot_solution"))

    # -- ABI marshalling -------------------------------------------------------------------------------
    def abi_sets(self, points):
        """The reference's point tuple -> point sets in ABI order."""
        return list(points)

    def abi_params(self, params):
        raise NotImplementedError

    def loss_fn(self, field_fn, points, params):
        if not cfield.is_siren_model(field_fn):
            raise TypeError("metapde_b200: loss_fn needs a SIREN Model (compat.field.Model); arbitrary field callables "
                            "cannot be fused and there is no fallback path")
        K = ops.kernels(field_fn.theta.device)
        sets = [torch.as_tensor(s, dtype=torch.float32, device=field_fn.theta.device).reshape(-1, 2).contiguous()
                for s in self.abi_sets(points)]
        prm = self.abi_params(params).to(field_fn.theta.device).reshape(1, -1).contiguous()
        call = _Call(K, field_fn.net_desc(K), self.name, sets, prm)
        terms = _TermsFn.apply(field_fn.theta, call)
        names = PDE_TERMS[self.name]
        b = {n: terms[names.index(n)] for n in _BOUNDARY_TERMS[self.name]}
        d = {n: terms[names.index(n)] for n in _DOMAIN_TERMS[self.name]}
        return b, d


def _fixed_key(key):
    if FLAGS.fixed_num_pdes is not None:
        return jr.PRNGKey(FLAGS.seed)
    return key


class PoissonDef(PdeDef):
    """src/poisson/poisson_def.py, poisson_common.py"""
    name, dim = "poisson", 1
    BaseField = cfield.NeuralPotential

    def sample_params(self, key):                                   # poisson_common.py:42-66
        k1, k2, k3 = jr.split(_fixed_key(_np(key)), 3)
        k1, k2, k3 = k1 * np.uint32(FLAGS.vary_source), k2 * np.uint32(FLAGS.vary_bc), k3 * np.uint32(FLAGS.vary_geometry)
        source = jr.normal(k1, (2, 3))
        bc = np.float32(FLAGS.bc_scale) * jr.uniform(k2, (5,), minval=-1.0, maxval=1.0)
        geo = jr.uniform(k3, (2,), minval=-0.2, maxval=0.2)
        return _to_dev(source), _to_dev(bc), _to_dev(geo)

    def sample_points(self, key, n, params):                        # :69-73
        k1, k2 = jr.split(_np(key))
        return (self.sample_points_on_boundary(k1, n, params), self.sample_points_in_domain(k2, n, params))

    def sample_points_on_boundary(self, key, n, params):            # :76-88
        c1, c2 = (float(v) for v in _np(params[2]))
        theta = np.linspace(0.0, 2 * np.pi, n, dtype=np.float32) + jr.uniform(_np(key), (n,), minval=0.0, maxval=2 * np.pi / n)
        r0 = 1.0 + c1 * np.cos(4 * theta) + c2 * np.cos(8 * theta)
        return _to_dev(np.stack([r0 * np.cos(theta), r0 * np.sin(theta)], axis=1))

    def sample_points_in_domain(self, key, n, params):              # :91-119 (incl. the arctan2(x, y) quirk)
        c1, c2 = (float(v) for v in _np(params[2]))
        k1, k2, k3 = jr.split(_np(key), 3)
        xs = jr.uniform(k1, (3 * n,), minval=FLAGS.xmin, maxval=FLAGS.xmax)
        ys = jr.uniform(k2, (3 * n,), minval=FLAGS.ymin, maxval=FLAGS.ymax)
        theta = np.arctan2(xs, ys)
        r0 = 1.0 + c1 * np.cos(4 * theta) + c2 * np.cos(8 * theta)
        in_hole = r0 < np.sqrt(xs * xs + ys * ys) + 1e-7
        idxs = jr.choice_without_replacement(k3, 3 * n, n, 1.0 - in_hole.astype(np.float32))
        return _to_dev(np.stack([xs, ys], axis=1)[idxs])

    def abi_params(self, params):
        s, b, g = params
        return torch.cat([torch.as_tensor(s, dtype=torch.float32).reshape(-1), torch.as_tensor(b, dtype=torch.float32).reshape(-1),
                          torch.as_tensor(g, dtype=torch.float32).reshape(-1)])


class TdBurgersDef(PdeDef):
    """src/burgers/td_burgers_def.py, td_burgers_common.py, burgers_formulation/default.py"""
    name, dim = "td_burgers", 1
    BaseField = cfield.NeuralPotential

    def sample_params(self, key):                                   # td_burgers_common.py:94-114
        k1, k2, _ = jr.split(_fixed_key(_np(key)), 3)
        k1, k2 = k1 * np.uint32(FLAGS.vary_source), k2 * np.uint32(FLAGS.vary_ic)
        source = np.float32(FLAGS.max_reynolds) * jr.uniform(k1, (1,), minval=0.8, maxval=1.0)
        ic = jr.uniform(k2, (2,), minval=-2.0, maxval=2.0)
        return _to_dev(source), _to_dev(ic)

    def _sample_time(self, key, n):                                 # :177-186
        nt = FLAGS.sample_tsteps - 1
        if FLAGS.sample_time_random:
            return jr.uniform(key, (nt * n, 1), minval=FLAGS.tmin, maxval=FLAGS.tmax)
        t = np.linspace(FLAGS.tmin, FLAGS.tmax, nt, endpoint=False, dtype=np.float32)
        return np.repeat(t[1:], n).reshape(-1, 1)

    def _on_side(self, key, x_val):                                 # :133-155
        _, k2 = jr.split(key)
        t = self._sample_time(k2, 1)
        x = np.full((t.shape[0], 1), x_val, np.float32)
        return np.concatenate([x, t], axis=1)

    def _in_domain(self, key, n):                                   # :158-169
        k1, k2 = jr.split(key, 2)
        n_scaled = n // (FLAGS.sample_tsteps - 1)
        n_sample = n_scaled * (FLAGS.sample_tsteps - 1)
        xs = jr.uniform(k1, (n_sample,), minval=FLAGS.xmin, maxval=FLAGS.xmax)
        return np.concatenate([xs[:, None], self._sample_time(k2, n_scaled)], axis=1)

    def sample_points(self, key, n, params):                        # :117-130 (left and right share k2: quirk 7)
        k1, k2, k3, k4, k5 = jr.split(_np(key), 5)
        dom_i = self._in_domain(k3, n)
        xi = np.concatenate([dom_i[:, 0:1], np.array([[FLAGS.xmin], [FLAGS.xmax]], np.float32)], axis=0)
        initial = np.concatenate([xi, np.zeros_like(xi)], axis=1)  # :172-176
        return (_to_dev(self._on_side(k2, FLAGS.xmin)), _to_dev(self._on_side(k2, FLAGS.xmax)), _to_dev(initial),
                _to_dev(self._in_domain(k5, n)))

    def sample_points_in_domain(self, key, n, params):
        return _to_dev(self._in_domain(_np(key), n))

    def abi_params(self, params):
        s, ic = params
        return torch.cat([torch.as_tensor(s, dtype=torch.float32).reshape(-1), torch.as_tensor(ic, dtype=torch.float32).reshape(-1)])


class HyperElasticityDef(PdeDef):
    """src/elasticity/hyper_elasticity_def.py, hyper_elasticity_common.py"""
    name, dim = "hyper_elasticity", 3                                # hyper_elasticity_def.py:18 (quirk 10)
    BaseField = cfield.make_nf_ndim(2)

    def sample_params(self, key):                                   # hyper_elasticity_common.py:89-180
        key, _ = jr.split(_fixed_key(_np(key)))
        H = FLAGS.max_holes
        if not H > 0:
            raise NotImplementedError("max_holes <= 0 breaks the reference's sample_params (SURVEY.md quirk 10)")
        while True:
            key, subkey = jr.split(key)
            ks = jr.split(key, 7)
            vs, vb, vg = np.uint32(FLAGS.vary_source), np.uint32(FLAGS.vary_bc), np.uint32(FLAGS.vary_geometry)
            source = jr.uniform(ks[0] * vs, (2,), minval=0.25, maxval=0.75)
            bc = np.float32(FLAGS.bc_scale) * jr.uniform(ks[1] * vb, (2,), minval=0.9, maxval=1.1)
            n_holes = H * H
            shape = 0.0 * np.array([jr.uniform(ks[2] * vg, (1,), minval=-0.1, maxval=0.1),
                                    jr.uniform(ks[3] * vg, (1,), minval=-0.1, maxval=0.1)], np.float32).T
            shapes = np.repeat(shape, n_holes, axis=0)
            x0 = np.linspace(FLAGS.xmin, FLAGS.xmax, H, dtype=np.float32)
            y0 = np.linspace(FLAGS.ymin, FLAGS.ymax, H, dtype=np.float32)
            x0y0 = np.array([(a, b) for a in x0 for b in y0], np.float32)
            L0 = x0[1] - x0[0]
            r0 = L0 * np.sqrt(2 * 0.5) / np.sqrt((2 + shapes[0, 0] ** 2 + shapes[0, 1] ** 2) * np.pi)
            scale = jr.uniform(ks[5] * vg, (1,), minval=0.2 * FLAGS.max_hole_size, maxval=1.5 * FLAGS.max_hole_size)
            sizes = np.full((n_holes, 1), r0, np.float32) * scale[0]
            theta = np.linspace(0, 2 * np.pi, 1000, dtype=np.float32)
            r_theta = scale * r0 * (1 + shape[0, 0] * np.cos(4 * theta) + shape[0, 1] * np.cos(8 * theta))
            tmin = (L0 - 2 * np.max(r_theta * np.cos(theta))) / L0
            per_hole = np.concatenate([shapes, x0y0, sizes], axis=1).astype(np.float32)
            key = subkey
            if not (tmin < 0.05):          # the while_loop continues while `feasibility` (= tmin < t_bar) holds
                break
        return _to_dev(source), _to_dev(bc), _to_dev(per_hole), n_holes

    @staticmethod
    def _in_hole(xy, per_hole, n_holes):                            # :183-189,206-216
        v = xy[:, None, :] - per_hole[None, :, 2:4]
        theta = np.arctan2(v[..., 0], v[..., 1])
        r0 = per_hole[None, :, 4] * (1.0 + per_hole[None, :, 0] * np.cos(4 * theta) + per_hole[None, :, 1] * np.cos(8 * theta))
        inh = r0 > np.linalg.norm(v, axis=-1) + 1e-7
        inh = inh & (np.arange(per_hole.shape[0])[None, :] < n_holes)
        return inh.any(axis=1)

    def _mask(self, key, n, xy, params):                            # :203-221
        inh = self._in_hole(xy, _np(params[2]), int(params[3]))
        return xy[jr.choice_without_replacement(key, xy.shape[0], n, 1.0 - inh.astype(np.float32))]

    def _edge(self, key, n, params, axis, val):                     # :224-283 (uniform draws use `key`, the mask k2)
        _, k2 = jr.split(key)
        lo, hi = (FLAGS.xmin, FLAGS.xmax) if axis == 0 else (FLAGS.ymin, FLAGS.ymax)
        u = jr.uniform(key, (10 * n,), minval=lo, maxval=hi)
        c = np.full(10 * n, val, np.float32)
        xy = np.stack([u, c], axis=1) if axis == 0 else np.stack([c, u], axis=1)
        return self._mask(k2, n, xy, params)

    def _on_pores(self, key, n, params):                            # :290-320
        per_hole = _np(params[2])
        n_tmp = int(1.5 * n)
        th = jr.uniform(key, (n_tmp,), minval=0.0, maxval=1.0) * np.float32(2 * np.pi)
        xs, ys = [], []
        for c1, c2, x0, y0, size in per_hole:
            r0 = size * (1.0 + c1 * np.cos(4 * th) + c2 * np.cos(8 * th))
            xs.append(x0 + r0 * np.cos(th)); ys.append(y0 + r0 * np.sin(th))
        xy = np.stack([np.concatenate(xs), np.concatenate(ys)], axis=1)
        inb = (xy[:, 0] > FLAGS.xmin) & (xy[:, 0] < FLAGS.xmax) & (xy[:, 1] > FLAGS.ymin) & (xy[:, 1] < FLAGS.ymax)
        return xy[jr.choice_without_replacement(key, xy.shape[0], n, inb.astype(np.float32))]

    def _in_domain(self, key, n, params):                           # :323-347
        k1, k2, k3 = jr.split(key, 3)
        xs = jr.uniform(k1, (3 * n,), minval=FLAGS.xmin, maxval=FLAGS.xmax)
        ys = jr.uniform(k2, (3 * n,), minval=FLAGS.ymin, maxval=FLAGS.ymax)
        return self._mask(k3, n, np.stack([xs, ys], axis=1), params)

    def sample_points(self, key, n, params):                        # :192-200
        k1, k2, k3, k4, k5, k6 = jr.split(_np(key), 6)
        top = self._edge(k1, n, params, 0, FLAGS.ymax)
        bottom = self._edge(k2, n, params, 0, FLAGS.ymin)
        left = self._edge(k3, n, params, 1, FLAGS.xmin)
        right = self._edge(k4, n, params, 1, FLAGS.xmax)
        holes = self._on_pores(k5, n, params) if FLAGS.max_holes > 0 else top
        return tuple(_to_dev(a) for a in (top, bottom, left, right, holes, self._in_domain(k6, n, params)))

    def sample_points_in_domain(self, key, n, params):
        return _to_dev(self._in_domain(_np(key), n, params))

    def abi_sets(self, points):
        return [points[0], points[1], points[5]]                    # top, bottom, domain enter the loss (:69-86)

    def abi_params(self, params):
        return torch.cat([torch.as_tensor(params[0], dtype=torch.float32).reshape(-1),
                          torch.as_tensor(params[1], dtype=torch.float32).reshape(-1)])


poisson_def, td_burgers_def, hyper_elasticity_def = PoissonDef(), TdBurgersDef(), HyperElasticityDef()


def get_pde(pde_name):
    """src/get_pde.py:10-24 (including the FLAGS.domain_loss side effect)."""
    if pde_name == "poisson":
        FLAGS.domain_loss = "domain_loss"
        return poisson_def
    elif pde_name == "td_burgers":
        FLAGS.domain_loss = "loss_domain,loss_initial"   # the reference assigns a list to this string flag (quirk 13)
        return td_burgers_def
    elif pde_name == "hyper_elasticity":
        FLAGS.domain_loss = "loss_domain"
        return hyper_elasticity_def
    else:
        raise Exception("Invalid PDE {}".format(pde_name))
