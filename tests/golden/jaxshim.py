"""A torch-backed stand-in for the ``jax`` / ``flax`` (0.2.20 / 0.3.3 era) modules -- FIXTURE GENERATION ONLY.

Purpose: execute the reference's OWN source files for the hot path
(/root/reference/src/nets/{field,maml,leap}.py, src/poisson/poisson_common.py,
src/burgers/td_burgers_common.py + burgers_formulation/default.py,
src/elasticity/hyper_elasticity_common.py) unmodified, in this container, where JAX
and Flax cannot be installed (no wheel, no network), and record their outputs as golden
fixtures (tests/golden/make_reference_fixtures.py).  Only the third-party API those
files call is provided here, restated from the published semantics of those libraries:

  jax.numpy        -> torch (float64: the shim is an "x64" run of the reference math)
  jax.vmap         -> a Python loop over the mapped axis + stack (supports data-dependent lax.cond)
  jax.jacfwd/jacrev/grad/value_and_grad -> torch.func / torch.autograd (create_graph, so they nest)
  jax.lax.scan/cond, jax.tree_util.*, jax.random.{PRNGKey,split,uniform,normal}
  flax.nn.Module (pre-Linen: apply / partial / init_by_shape / param / auto-named submodules),
  flax.nn.Dense, flax.nn.Model, flax.nn.initializers, flax.optim.{Momentum,Adam}

The PRNG is NOT threefry (keys are hashed into numpy generators): the fixtures store the
arrays the reference code saw, so bit-parity with jax.random is not needed.  fenics /
matplotlib / mshr are replaced by inert stubs (plots and FEM are out of the path).

This file is test infrastructure; nothing under meta_pde_b200/ imports it.
"""
import math
import sys
import types
from collections import namedtuple

import numpy as _onp
import torch
import torch.nn.functional as _F

DT = torch.float64


# ------------------------------------------------------------------------------------------------ tensor patches
def _patch_tensor():
    if getattr(torch.Tensor, "_jaxshim_patched", False):
        return
    _orig_getitem = torch.Tensor.__getitem__
    _orig_transpose = torch.Tensor.transpose

    def _getitem(self, idx):
        if isinstance(idx, slice) and idx.step == -1 and idx.start is None and idx.stop is None:
            return torch.flip(self, [0])          # numpy's x[::-1]
        return _orig_getitem(self, idx)

    def _transpose(self, *a, **k):
        if not a and not k:
            return self.T                          # numpy's x.transpose()
        return _orig_transpose(self, *a, **k)

    torch.Tensor.__getitem__ = _getitem
    torch.Tensor.transpose = _transpose
    torch.Tensor.astype = lambda self, dt: self.to(_dtype(dt))
    _orig_rsub = torch.Tensor.__rsub__
    torch.Tensor.__rsub__ = lambda self, o: _orig_rsub(self.to(torch.int64) if self.dtype == torch.bool else self, o)   # numpy's 1 - mask
    torch.Tensor._jaxshim_patched = True


def _dtype(dt):
    if dt in (float, "float32", "float64", _onp.float32, _onp.float64, torch.float32, torch.float64, None):
        return DT
    if dt in (int, _onp.int32, _onp.int64, torch.int32, torch.int64):
        return torch.int64
    if dt is bool:
        return torch.bool
    return DT


def _t(x):
    if torch.is_tensor(x):
        return x
    if isinstance(x, (list, tuple)) and any(torch.is_tensor(e) for e in x):
        return torch.stack([_t(e) for e in x])
    a = _onp.asarray(x)
    if a.dtype.kind == "f":
        return torch.as_tensor(a, dtype=DT)
    return torch.as_tensor(a)


# ------------------------------------------------------------------------------------------------ pytrees
class _Leaf:
    def __repr__(self):
        return "*"

    def __eq__(self, o):
        return isinstance(o, _Leaf)

    def __hash__(self):
        return 1


LEAF = _Leaf()
_REGISTRY = {}   # type -> (flatten -> (children, aux), unflatten(aux, children))


def register_pytree_node(cls, flatten, unflatten):
    _REGISTRY[cls] = (flatten, unflatten)


def tree_flatten(tree):
    leaves = []

    def rec(t):
        if t is None:
            return ("none",)
        if type(t) in _REGISTRY:
            children, aux = _REGISTRY[type(t)][0](t)
            return ("custom", type(t), aux, tuple(rec(c) for c in children))
        if isinstance(t, tuple) and hasattr(t, "_fields"):
            return ("namedtuple", type(t), tuple(rec(c) for c in t))
        if isinstance(t, tuple):
            return ("tuple", tuple(rec(c) for c in t))
        if isinstance(t, list):
            return ("list", tuple(rec(c) for c in t))
        if isinstance(t, dict):
            keys = sorted(t.keys())
            return ("dict", tuple(keys), tuple(rec(t[k]) for k in keys))
        leaves.append(t)
        return LEAF

    return leaves, rec(tree)


def tree_unflatten(treedef, leaves):
    it = iter(leaves)

    def rec(d):
        if d == LEAF:
            return next(it)
        kind = d[0]
        if kind == "none":
            return None
        if kind == "custom":
            return _REGISTRY[d[1]][1](d[2], [rec(c) for c in d[3]])
        if kind == "namedtuple":
            return d[1](*[rec(c) for c in d[2]])
        if kind == "tuple":
            return tuple(rec(c) for c in d[1])
        if kind == "list":
            return [rec(c) for c in d[1]]
        if kind == "dict":
            return {k: rec(c) for k, c in zip(d[1], d[2])}
        raise TypeError(d)

    return rec(treedef)


def tree_leaves(tree):
    return tree_flatten(tree)[0]


def tree_structure(tree):
    return tree_flatten(tree)[1]


def tree_map(f, tree, *rest):
    leaves, td = tree_flatten(tree)
    others = [tree_flatten(r)[0] for r in rest]
    for o in others:
        assert len(o) == len(leaves), "tree_multimap: mismatching trees"
    return tree_unflatten(td, [f(*xs) for xs in zip(leaves, *others)])


def tree_reduce(f, tree):
    leaves = tree_leaves(tree)
    acc = leaves[0]
    for x in leaves[1:]:
        acc = f(acc, x)
    return acc


# ------------------------------------------------------------------------------------------------ transformations
def jit(fn=None, static_argnums=None, **kw):
    if fn is None:
        return lambda f: f
    return fn


def vmap(fn, in_axes=0, out_axes=0):
    def wrapped(*args):
        axes = tuple(in_axes) if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = None
        for a, ax in zip(args, axes):
            if ax is not None:
                n = tree_leaves(a)[0].shape[ax]
                break
        outs = []
        for i in range(n):
            sl = [tree_map(lambda x: torch.select(_t(x), ax, i), a) if ax is not None else a for a, ax in zip(args, axes)]
            outs.append(fn(*sl))
        return tree_map(lambda *xs: torch.stack([_t(x).to(DT) if _t(x).is_floating_point() else _t(x) for x in xs], dim=out_axes),
                        outs[0], *outs[1:])

    return wrapped


def jacfwd(fn, argnums=0):
    return torch.func.jacfwd(lambda *a: _t(fn(*a)), argnums=argnums)


def jacrev(fn, argnums=0):
    return torch.func.jacrev(lambda *a: _t(fn(*a)), argnums=argnums)


def value_and_grad(fn, argnums=0, has_aux=False):
    def wrapped(*args):
        arg = args[argnums]
        leaves, td = tree_flatten(arg)
        # leaves that are already part of an outer differentiation keep their graph (nested grad);
        # free leaves become fresh differentiable inputs
        live = [l if (torch.is_tensor(l) and l.requires_grad) else _t(l).detach().clone().requires_grad_(True) for l in leaves]
        new_args = list(args)
        new_args[argnums] = tree_unflatten(td, live)
        out = fn(*new_args)
        val, aux = out if has_aux else (out, None)
        grads = torch.autograd.grad(val, live, create_graph=True, allow_unused=True)
        grads = [g if g is not None else torch.zeros_like(l) for g, l in zip(grads, live)]
        g = tree_unflatten(td, grads)
        return ((val, aux), g) if has_aux else (val, g)

    return wrapped


def grad(fn, argnums=0, has_aux=False):
    vg = value_and_grad(fn, argnums, has_aux)

    def wrapped(*args):
        v, g = vg(*args)
        return (g, v[1]) if has_aux else g

    return wrapped


def _scan(f, init, xs, length=None):
    n = length if xs is None else tree_leaves(xs)[0].shape[0]
    carry, ys = init, []
    for i in range(n):
        x = None if xs is None else tree_map(lambda a: a[i], xs)
        carry, y = f(carry, x)
        ys.append(y)
    stacked = tree_map(lambda *a: torch.stack([_t(e) for e in a]), ys[0], *ys[1:])
    return carry, stacked


def _cond(pred, true_fn, false_fn, operand):
    return true_fn(operand) if bool(pred) else false_fn(operand)


# ------------------------------------------------------------------------------------------------ random
def _gen(key, salt):
    k = [int(x) & 0xFFFFFFFF for x in _t(key).reshape(-1).tolist()]
    return _onp.random.default_rng(k + [salt])


def PRNGKey(seed):
    return torch.tensor([0, int(seed) & 0xFFFFFFFF], dtype=torch.int64)


def split(key, num=2):
    return torch.as_tensor(_gen(key, 0x51).integers(0, 2 ** 32, size=(int(num), 2)), dtype=torch.int64)


def uniform(key, shape=(), dtype=None, minval=0.0, maxval=1.0):
    u = _gen(key, 0x52).random(size=tuple(shape))
    return torch.as_tensor(u, dtype=DT) * (float(maxval) - float(minval)) + float(minval)


def normal(key, shape=(), dtype=None):
    return torch.as_tensor(_gen(key, 0x53).standard_normal(size=tuple(shape)), dtype=DT)


def use_threefry(jr):
    """Swap the stand-in's hashed PRNG for `jr` (meta_pde_b200.compat.random: the Threefry-2x32 restatement of jax 0.2.20's
    jax.random) so that the reference's SAMPLER and INIT source can be executed with the key stream the compat layer uses.
    Keys are int64 tensors holding the two uint32 words (so that the reference's `key * FLAGS.vary_x` works)."""
    def k_(key):
        return _onp.asarray([int(v) & 0xFFFFFFFF for v in _t(key).reshape(-1).tolist()], dtype=_onp.uint32)

    def f_(x):
        return float(x.item()) if torch.is_tensor(x) else float(x)

    def out_key(a):
        return torch.as_tensor(_onp.asarray(a, dtype=_onp.int64))

    def t_uniform(key, shape=(), dtype=None, minval=0.0, maxval=1.0):
        return torch.as_tensor(jr.uniform(k_(key), tuple(shape), minval=f_(minval), maxval=f_(maxval)).astype(_onp.float64))

    def t_choice(key, a, shape=(), replace=True, p=None):
        assert not replace and p is not None and isinstance(a, int)
        idx = jr.choice_without_replacement(k_(key), int(a), int(shape[0]), _t(p).to(torch.float32).numpy())
        return torch.as_tensor(_onp.asarray(idx, dtype=_onp.int64))

    rnd = sys.modules["jax.random"]
    rnd.PRNGKey = lambda seed: out_key(jr.PRNGKey(int(seed)))
    rnd.split = lambda key, num=2: out_key(jr.split(k_(key), int(num)))
    rnd.uniform = t_uniform
    rnd.normal = lambda key, shape=(), dtype=None: torch.as_tensor(jr.normal(k_(key), tuple(shape)).astype(_onp.float64))
    rnd.choice = t_choice
    rnd.randint = lambda key, shape, minval, maxval: torch.as_tensor(
        jr.randint(k_(key), tuple(shape), int(_t(minval).reshape(-1)[0]), int(_t(maxval).reshape(-1)[0])).astype(_onp.int64))


def _while_loop(cond, body, init):
    val = init
    while bool(cond(val)):
        val = body(val)
    return val


# ------------------------------------------------------------------------------------------------ flax.nn
_stack = []


class _Frame:
    def __init__(self, params, rng, init):
        self.params, self.rng, self.init, self.counters = params, rng, init, {}


def _fold(rng, name):
    if rng is None:
        return None
    h = sum((i + 1) * ord(c) for i, c in enumerate(name)) & 0xFFFFFFFF
    return torch.as_tensor(_gen(rng, h).integers(0, 2 ** 32, size=(2,)), dtype=torch.int64)


class Module:
    """flax.nn.Module (pre-Linen): subclasses define apply(self, *args, **kwargs)."""
    _defaults = {}

    def __new__(cls, *args, name=None, **kwargs):
        if not _stack:
            raise RuntimeError("flax shim: a Module can only be applied inside init/call")
        parent = _stack[-1]
        base = cls._base_name()
        if name is None:
            idx = parent.counters.get(base, 0)
            parent.counters[base] = idx + 1
            name = f"{base}_{idx}"
        if parent.init:
            child_params = parent.params.setdefault(name, {})
        else:
            child_params = parent.params[name]
        return cls._run(_Frame(child_params, _fold(parent.rng, name), parent.init), args, kwargs)

    @classmethod
    def _base_name(cls):
        c = cls
        while getattr(c, "_is_partial", False):
            c = c.__bases__[0]
        return c.__name__

    @classmethod
    def _run(cls, frame, args, kwargs):
        inst = object.__new__(cls)
        inst._frame = frame
        _stack.append(frame)
        try:
            return inst.apply(*args, **{**cls._defaults, **kwargs})
        finally:
            _stack.pop()

    @classmethod
    def partial(cls, **kwargs):
        return type(cls.__name__, (cls,), {"_defaults": {**cls._defaults, **kwargs}, "_is_partial": True})

    @classmethod
    def init(cls, rng, *args, **kwargs):
        frame = _Frame({}, rng, True)
        out = cls._run(frame, args, kwargs)
        return out, frame.params

    @classmethod
    def init_by_shape(cls, rng, input_specs, **kwargs):
        inputs = [torch.zeros(tuple(shape), dtype=DT) for shape, _ in input_specs]
        return cls.init(rng, *inputs, **kwargs)

    @classmethod
    def call(cls, params, *args, **kwargs):
        return cls._run(_Frame(params, None, False), args, kwargs)

    def param(self, name, shape, initializer):
        fr = self._frame
        if fr.init:
            if name in fr.params:
                raise ValueError(f"flax shim: duplicate param {name}")
            fr.params[name] = _t(initializer(_fold(fr.rng, name), tuple(shape))).to(DT)
        return fr.params[name]


def _zeros_init(key, shape, dtype=None):
    return torch.zeros(tuple(shape), dtype=DT)


def _ones_init(key, shape, dtype=None):
    return torch.ones(tuple(shape), dtype=DT)


def _lecun_normal():
    def init(key, shape, dtype=None):
        return normal(key, shape) / math.sqrt(shape[0])
    return init


class Dense(Module):
    """flax.nn.Dense: y = x @ kernel + bias, kernel [n_in, features]; params created kernel first."""

    def apply(self, inputs, features, use_bias=True, kernel_init=None, bias_init=_zeros_init, **unused):
        kernel_init = kernel_init or _lecun_normal()
        inputs = _t(inputs)
        kernel = self.param("kernel", (inputs.shape[-1], features), kernel_init)
        y = inputs @ kernel
        if use_bias:
            y = y + self.param("bias", (features,), bias_init)
        return y


class Model:
    """flax.nn.Model(module, params): a callable pytree whose leaves are the params."""

    def __init__(self, module, params):
        self.module, self.params = module, params

    def __call__(self, *args, **kwargs):
        return self.module.call(self.params, *args, **kwargs)

    def replace(self, **kw):
        return Model(kw.get("module", self.module), kw.get("params", self.params))


register_pytree_node(Model, lambda m: ((m.params,), m.module), lambda aux, ch: Model(aux, ch[0]))


# ------------------------------------------------------------------------------------------------ flax.optim
class Optimizer:
    def __init__(self, optimizer_def, state, target):
        self.optimizer_def, self.state, self.target = optimizer_def, state, target

    def apply_gradient(self, grads, **hyper):
        target, state = self.optimizer_def.apply_gradient(self.target, self.state, grads, **hyper)
        return Optimizer(self.optimizer_def, state, target)


register_pytree_node(Optimizer, lambda o: ((o.target, o.state), o.optimizer_def),
                     lambda aux, ch: Optimizer(aux, ch[1], ch[0]))


class Momentum:
    """flax 0.3.3 optim.Momentum (no nesterov, no weight decay): m <- beta m + g ; p <- p - lr m."""

    def __init__(self, learning_rate=None, beta=0.9, weight_decay=0, nesterov=False):
        assert not nesterov and not weight_decay
        self.lr, self.beta = learning_rate, beta

    def create(self, target):
        return Optimizer(self, {"step": 0, "momentum": tree_map(lambda p: torch.zeros_like(_t(p)), target)}, target)

    def apply_gradient(self, target, state, grads):
        mom = tree_map(lambda m, g: self.beta * m + g, state["momentum"], grads)
        new_t = tree_map(lambda p, m: p - self.lr * m, target, mom)
        return new_t, {"step": state["step"] + 1, "momentum": mom}


class Adam:
    """flax 0.3.3 optim.Adam: bias-corrected moments, denom = sqrt(v_hat) + eps, no weight decay."""

    def __init__(self, learning_rate=None, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0):
        assert not weight_decay
        self.lr, self.b1, self.b2, self.eps = learning_rate, beta1, beta2, eps

    def create(self, target):
        z = lambda: tree_map(lambda p: torch.zeros_like(_t(p)), target)
        return Optimizer(self, {"step": 0, "m": z(), "v": z()}, target)

    def apply_gradient(self, target, state, grads):
        t = state["step"] + 1
        m = tree_map(lambda m_, g: self.b1 * m_ + (1.0 - self.b1) * g, state["m"], grads)
        v = tree_map(lambda v_, g: self.b2 * v_ + (1.0 - self.b2) * g * g, state["v"], grads)
        new_t = tree_map(lambda p, m_, v_: p - self.lr * (m_ / (1.0 - self.b1 ** t)) /
                         (torch.sqrt(v_ / (1.0 - self.b2 ** t)) + self.eps), target, m, v)
        return new_t, {"step": t, "m": m, "v": v}


def jax_adam(step_size, b1=0.9, b2=0.999, eps=1e-8):
    """jax.experimental.optimizers.adam -> (init, update, get_params); x - lr * mhat / (sqrt(vhat) + eps)."""
    def init(x0):
        return (x0, tree_map(lambda p: torch.zeros_like(_t(p)), x0), tree_map(lambda p: torch.zeros_like(_t(p)), x0))

    def update(i, g, state):
        x, m, v = state
        m = tree_map(lambda m_, g_: (1 - b1) * g_ + b1 * m_, m, g)
        v = tree_map(lambda v_, g_: (1 - b2) * g_ * g_ + b2 * v_, v, g)
        x = tree_map(lambda x_, m_, v_: x_ - step_size * (m_ / (1 - b1 ** (i + 1))) /
                     (torch.sqrt(v_ / (1 - b2 ** (i + 1))) + eps), x, m, v)
        return (x, m, v)

    return init, update, lambda state: state[0]


# ------------------------------------------------------------------------------------------------ jax.numpy
def _make_jnp():
    m = types.ModuleType("jax.numpy")
    m.pi = math.pi
    m.float32, m.float64, m.float, m.int32, m.int = _onp.float32, _onp.float64, float, _onp.int32, int
    u = lambda f: (lambda x, *a, **k: f(_t(x), *a, **k))
    for name in ["sin", "cos", "exp", "log", "sqrt", "abs", "tanh", "arctan", "square"]:
        setattr(m, name, u(getattr(torch, {"arctan": "atan"}.get(name, name))))
    m.arctan2 = lambda a, b: torch.atan2(_t(a), _t(b))

    def _ax(k):
        if "axis" in k:
            k = dict(k)
            k["dim"] = k.pop("axis")
        return k

    m.mean = lambda x, *a, **k: torch.mean(_t(x), *a, **_ax(k))
    m.sum = lambda x, *a, **k: torch.sum(_t(x), *a, **_ax(k))
    m.max = lambda x, *a, **k: torch.max(_t(x), *a, **_ax(k))
    m.any = lambda x, *a, **k: torch.any(_t(x), *a, **_ax(k))
    m.trace = lambda x: torch.trace(_t(x))
    m.stack = lambda xs, axis=0: torch.stack([_t(x) for x in xs], dim=axis)
    m.concatenate = lambda xs, axis=0: torch.cat([_t(x) for x in xs], dim=axis)
    m.append = lambda a, b: torch.cat([_t(a).reshape(-1), _t(b).reshape(-1).to(_t(a).dtype)])
    m.array = lambda x, dtype=None: _t(x)
    m.asarray = m.array
    m.ones = lambda shape, dtype=None: torch.ones(tuple(shape) if not isinstance(shape, int) else (shape,), dtype=_dtype(dtype))
    m.zeros = lambda shape, dtype=None: torch.zeros(tuple(shape) if not isinstance(shape, int) else (shape,), dtype=_dtype(dtype))
    m.ones_like = lambda x: torch.ones_like(_t(x))
    m.zeros_like = lambda x: torch.zeros_like(_t(x))
    m.identity = lambda n: torch.eye(n, dtype=DT)
    m.eye = m.identity
    def _linspace(a, b, n, endpoint=True):
        a, b, n = float(a), float(b), int(n)
        return torch.linspace(a, b, n, dtype=DT) if endpoint else torch.linspace(a, b, n + 1, dtype=DT)[:-1]
    m.linspace = _linspace
    m.arange = lambda *a, dtype=None: torch.arange(*a)
    def _repeat(x, n, axis=None):
        x = _t(x)
        if x.dim() == 0 or axis is None:
            return torch.repeat_interleave(x.reshape(-1), int(n))
        return torch.repeat_interleave(x, int(n), dim=axis)
    m.repeat = _repeat
    m.tile = lambda x, reps: _t(x).repeat(*reps)
    m.matmul = lambda a, b: _t(a) @ _t(b)
    m.dot = m.matmul
    m.transpose = lambda x: _t(x).T
    m.squeeze = lambda x: _t(x).squeeze()
    m.reshape = lambda x, s: _t(x).reshape(s)
    m.isclose = lambda a, b: torch.isclose(_t(a), _t(b))
    m.where = lambda c, a, b: torch.where(c, _t(a), _t(b))
    m.linalg = types.SimpleNamespace(det=lambda x: torch.linalg.det(_t(x)), norm=lambda x: torch.linalg.norm(_t(x)))
    m.ndarray = torch.Tensor
    return m


# ------------------------------------------------------------------------------------------------ inert stubs
class _Inert(types.ModuleType):
    """fenics / matplotlib / mshr / dolfin: any attribute is an inert callable object."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()


class _Anything:
    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()


def install():
    """Put the stand-in modules into sys.modules (refuses if a real jax is importable)."""
    _patch_tensor()
    if "jax" in sys.modules and not getattr(sys.modules["jax"], "_is_shim", False):
        raise RuntimeError("a real jax is loaded; the shim is only for images without it")
    jnp = _make_jnp()
    jax = types.ModuleType("jax")
    jax._is_shim = True
    jax.numpy = jnp
    jax.jit, jax.vmap, jax.grad, jax.value_and_grad, jax.jacfwd, jax.jacrev = jit, vmap, grad, value_and_grad, jacfwd, jacrev
    jax.hessian = lambda f: jacfwd(jacrev(f))
    jax.checkpoint = lambda f: f
    jax.tree_map = tree_map
    tu = types.ModuleType("jax.tree_util")
    tu.tree_map, tu.tree_multimap, tu.tree_reduce = tree_map, tree_map, tree_reduce
    tu.tree_flatten, tu.tree_unflatten, tu.tree_leaves, tu.tree_structure = tree_flatten, tree_unflatten, tree_leaves, tree_structure
    jax.tree_util = tu
    lax = types.ModuleType("jax.lax")
    lax.scan, lax.cond, lax.while_loop = _scan, _cond, _while_loop
    jax.lax = lax
    rnd = types.ModuleType("jax.random")
    rnd.PRNGKey, rnd.split, rnd.uniform, rnd.normal = PRNGKey, split, uniform, normal
    jax.random = rnd
    jnn = types.ModuleType("jax.nn")
    jnn.softplus = lambda x: _F.softplus(_t(x))
    jax.nn = jnn
    cfg = types.ModuleType("jax.config")
    cfg.config = types.SimpleNamespace(update=lambda *a, **k: None)
    jax.config = cfg
    exp = types.ModuleType("jax.experimental")
    opt = types.ModuleType("jax.experimental.optimizers")
    opt.adam = jax_adam
    exp.optimizers = opt
    jax.experimental = exp
    jax.ops = types.SimpleNamespace()

    flax = types.ModuleType("flax")
    nn = types.ModuleType("flax.nn")
    nn.Module, nn.Dense, nn.Model = Module, Dense, Model
    nn.relu = lambda x: torch.relu(_t(x))
    nn.swish = lambda x: _t(x) * torch.sigmoid(_t(x))
    nn.module = types.SimpleNamespace()
    nn.initializers = types.SimpleNamespace(ones=_ones_init, zeros=_zeros_init, lecun_normal=_lecun_normal,
                                            variance_scaling=lambda *a, **k: _lecun_normal())
    flax.nn = nn
    optim = types.ModuleType("flax.optim")
    optim.Momentum, optim.Adam, optim.Optimizer = Momentum, Adam, Optimizer
    flax.optim = optim
    flax.serialization = types.SimpleNamespace()

    mods = {"jax": jax, "jax.numpy": jnp, "jax.tree_util": tu, "jax.lax": lax, "jax.random": rnd, "jax.nn": jnn,
            "jax.config": cfg, "jax.experimental": exp, "jax.experimental.optimizers": opt,
            "flax": flax, "flax.nn": nn, "flax.optim": optim}
    for name in ["fenics", "mshr", "dolfin", "matplotlib", "matplotlib.pyplot"]:
        mods[name] = _Inert(name)
    mods["matplotlib"].pyplot = mods["matplotlib.pyplot"]
    sys.modules.update(mods)
    return jax, flax
