"""SIREN neural field on the plugin surface (mirror of /root/reference/src/nets/field.py:23-44,140-226,259).

``NeuralPotential`` / ``make_nf_ndim(n)`` are Flax-0.3.3-style module classes: ``.partial(sizes=, dense_args=,
nonlinearity=, omega=, omega0=, log_scale=)`` binds constructor arguments, ``.init_by_shape(key, [((1, 2), f32)])``
returns ``(output, params)`` with the reference's parameter tree (``Dense_i/{kernel,bias}``, ``log_input_scale
[1, 2]``, ``log_output_scale [1, out]``), ``Model(Field, params)`` is the callable ``field_fn``.

Unlike Flax, a ``Model`` keeps ONE flat float32 CUDA tensor ``theta`` in the C-ABI order (include/metapde_b200.h);
``model.params`` exposes the reference's tree as views into it, so the fused kernels take the model without
copies.  ``model(x)`` runs ``mpde_field_eval``.  Only the SIREN configuration the kernels implement is accepted
(nonlinearity sin, in_dim 2, layer_size in {32, 64, 128, 256}); anything else raises, there is no fallback.
"""
import math

import numpy as np
import torch

from .. import ops
from . import random as jrandom
from .flags import FLAGS


def sin(x):
    """The ``np.sin`` the drivers pass as ``nonlinearity`` (identity marker on this surface)."""
    return torch.sin(x)


def swish(x):
    """The ``flax.nn.swish`` the drivers pass as ``nonlinearity`` under ``--nosiren`` (maml_pde.py:115)."""
    return x * torch.sigmoid(x)


def variance_scaling_fan_in_truncated_normal():
    """field.py:171-176 (non-sin branch): variance_scaling(1.0, "fan_in", "truncated_normal") =
    jax.random.truncated_normal(key, -2, 2) * sqrt(1 / fan_in) / 0.87962566103423978, the truncated normal drawn as
    sqrt(2) * erfinv(U(erf(-sqrt 2), erf(sqrt 2))) from the threefry uniform of compat.random (erf/erfinv by scipy in
    float64, rounded to float32: XLA's float32 erf_inv polynomial is not restated, so these draws are NOT pinned
    bit-for-bit -- the distribution and the key stream are)."""
    def init_fn(key, shape, dtype=np.float32):
        from scipy import special
        a, b = special.erf(-2.0 / np.sqrt(2.0)), special.erf(2.0 / np.sqrt(2.0))
        u = np.asarray(jrandom.uniform(key, shape, minval=np.float32(a), maxval=np.float32(b)), np.float64)
        z = np.clip(np.sqrt(2.0) * special.erfinv(u), np.nextafter(np.float32(-2.0), np.float32(0.0)),
                    np.nextafter(np.float32(2.0), np.float32(0.0)))
        return (z * np.sqrt(1.0 / shape[0]) / 0.87962566103423978).astype(np.float32)
    return init_fn


def siren_init(omega):
    """field.py:23-34: U(-sqrt(6 / fan_in) / omega, +sqrt(6 / fan_in) / omega) (threefry draws of compat.random)."""
    def init_fn(key, shape, dtype=np.float32, omega=omega):
        lim = np.sqrt(6.0 / shape[0]) / omega
        return jrandom.uniform(key, shape, minval=-lim, maxval=lim)
    return init_fn


def first_layer_siren_init(omega, omega0):
    """field.py:37-44: (omega0 / omega) * U(-1 / fan_in, +1 / fan_in)."""
    def init_fn(key, shape, dtype=np.float32, omega=omega, omega0=omega0):
        lim = 1.0 / shape[0]
        return np.float32(omega0 / omega) * jrandom.uniform(key, shape, minval=-lim, maxval=lim)
    return init_fn


def constant_init(val):
    """field.py:140-143: val * ones (the log io-scales)."""
    def initializer(key, shape, dtype=np.float32):
        return np.full(tuple(shape), val, np.float32)
    return initializer


class FieldSpec:
    """Bound constructor arguments of a field class."""

    def __init__(self, out_dim, sum_output, sizes=None, dense_args=(), nonlinearity=None, n_fourier=None,
                 log_scale=False, omega=30.0, omega0=30.0, use_laaf=False, use_nlaaf=False):
        self.out_dim, self.sum_output = out_dim, sum_output
        self.sizes = list(sizes) if sizes is not None else None
        self.nonlinearity, self.n_fourier = nonlinearity, n_fourier
        self.log_scale, self.omega, self.omega0 = bool(log_scale), float(omega), float(omega0)
        self.use_laaf, self.use_nlaaf = use_laaf, use_nlaaf

    def check(self):
        if self.sizes is None or len(self.sizes) < 1 or len(set(self.sizes)) != 1:
            raise NotImplementedError("metapde_b200: hidden layers must all have the same layer_size")
        if self.nonlinearity in (sin, np.sin, torch.sin):
            self.nonlinearity_name = "sin"
        elif self.nonlinearity in (swish, torch.nn.functional.silu):
            self.nonlinearity_name = "swish"
        else:
            raise NotImplementedError("metapde_b200: nonlinearity must be field.sin (np.sin) or field.swish (flax.nn.swish), "
                                      "the two the reference's drivers pass; there is no fallback path")
        if self.n_fourier is not None or self.use_laaf or self.use_nlaaf:
            raise NotImplementedError("metapde_b200: fourier features / LAAF are not implemented by the sm_100a kernels "
                                      "(no reference driver passes them: nn_pde.py:161-162 are commented out)")

    @property
    def L(self):
        return len(self.sizes)

    @property
    def W(self):
        return self.sizes[0]

    def param_shapes(self, in_dim=2):
        """[(path, shape)] in the C-ABI flat order."""
        out, n_in = [], in_dim
        for l, n_out in enumerate(self.sizes + [self.out_dim]):
            out.append(((f"Dense_{l}", "kernel"), (n_in, n_out)))
            out.append(((f"Dense_{l}", "bias"), (n_out,)))
            n_in = n_out
        if self.log_scale:
            out.append((("log_input_scale",), (1, in_dim)))
            out.append((("log_output_scale",), (1, self.out_dim)))
        return out

    def param_count(self):
        return sum(int(np.prod(s)) for _, s in self.param_shapes())


class _FieldMeta(type):
    def __repr__(cls):
        return f"<metapde_b200 field {cls.__name__} out_dim={cls._out_dim}>"


class _FieldBase(metaclass=_FieldMeta):
    _out_dim, _sum_output, _kwargs = 1, True, {}

    @classmethod
    def partial(cls, **kwargs):
        return _FieldMeta(cls.__name__, (cls,), {"_kwargs": {**cls._kwargs, **kwargs}})

    @classmethod
    def spec(cls):
        s = FieldSpec(cls._out_dim, cls._sum_output, **cls._kwargs)
        s.check()
        return s

    @classmethod
    def init_by_shape(cls, key, input_specs, **kwargs):
        """-> (output at zeros, params).  Distributions of field.py:23-44,140-143; draws from compat.random
        (threefry) with one subkey per parameter in creation order (Flax's own per-name key folding is not restated)."""
        spec = cls.partial(**kwargs).spec() if kwargs else cls.spec()
        (shape, _dtype), = input_specs
        if shape[-1] != 2:
            raise NotImplementedError("metapde_b200: in_dim must be 2 (src/maml_pde.py:124 hard-wires (1, 2))")
        shapes = spec.param_shapes()
        keys = jrandom.split(key, len(shapes))
        vals = []
        for (path, shp), k in zip(shapes, keys):
            if path[-1] == "kernel":
                if spec.nonlinearity_name == "swish":
                    init = variance_scaling_fan_in_truncated_normal()
                else:
                    init = first_layer_siren_init(spec.omega, spec.omega0) if path[0] == "Dense_0" else siren_init(spec.omega)
                v = init(k, shp)
            elif path[-1] == "bias":
                v = np.zeros(shp, np.float32)
            else:
                v = constant_init(math.log(1.0 / FLAGS.io_scale_lr_factor))(k, shp)
            vals.append(np.asarray(v, np.float32).reshape(-1))
        theta = torch.from_numpy(np.concatenate(vals))
        model = Model(cls, theta)
        out = None
        if torch.cuda.is_available():
            out = model(torch.zeros(*shape, device=model.theta.device))
        return out, model.params


class NeuralField1d(_FieldBase):
    """field.py:212-216: scalar field, the 1-wide output summed away."""
    _out_dim, _sum_output = 1, True


NeuralPotential = NeuralField1d   # field.py:259


def make_nf_ndim(n_dims):
    """field.py:219-226: vector-valued field."""
    return _FieldMeta("HigherDimNeuralField", (_FieldBase,), {"_out_dim": int(n_dims), "_sum_output": False})


def _views(theta, spec):
    """The reference's parameter tree as views into a flat theta (leading dims kept)."""
    tree, off = {}, 0
    lead = theta.shape[:-1]
    for path, shp in spec.param_shapes():
        n = int(np.prod(shp))
        v = theta[..., off:off + n].reshape(*lead, *shp)
        off += n
        if len(path) == 2:
            tree.setdefault(path[0], {})[path[1]] = v
        else:
            tree[path[0]] = v
    assert off == theta.shape[-1]
    return tree


def flatten_params(params, spec, device=None):
    """Reference parameter tree (torch / numpy leaves, optional leading dims) -> flat theta in ABI order."""
    parts = []
    for path, shp in spec.param_shapes():
        x = params[path[0]] if len(path) == 1 else params[path[0]][path[1]]
        x = torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x, dtype=torch.float32)
        lead = x.shape[:x.dim() - len(shp)]
        parts.append(x.reshape(*lead, -1))
    out = torch.cat(parts, dim=-1).contiguous()
    return out.to(device) if device is not None else out


class Model:
    """``flax.nn.Model(Field, params)``: the callable field, a pytree over one flat ``theta`` (any leading dims:
    an lr tree is ``Model(Field, theta[K, P])``)."""

    def __init__(self, module, params_or_theta, device=None):
        self.module = module
        self.spec = module.spec()
        if torch.is_tensor(params_or_theta):
            theta = params_or_theta
        else:
            theta = flatten_params(params_or_theta, self.spec)
        if theta.shape[-1] != self.spec.param_count():
            raise ValueError(f"expected {self.spec.param_count()} parameters, got {theta.shape[-1]}")
        if not theta.is_cuda and (device is not None or torch.cuda.is_available()):
            theta = theta.to(device or torch.device("cuda", torch.cuda.current_device()))
        # without CUDA the tree can still be built / (de)serialised; every compute entry point raises (ops.kernels)
        self.theta = theta.to(torch.float32)

    @property
    def params(self):
        return _views(self.theta, self.spec)

    def replace(self, params=None, theta=None):
        return Model(self.module, theta if theta is not None else params)

    def tree_map(self, fn, *others):
        return Model(self.module, fn(self.theta, *[o.theta for o in others]))

    def net_desc(self, K):
        s = self.spec
        return K.net_desc(s.L, s.W, s.out_dim, s.omega, s.log_scale, nonlinearity=s.nonlinearity_name)

    def __call__(self, x):
        """field_fn(x): x [N, 2] or [2] -> [N] (1-d field) / [N, out]."""
        K = ops.kernels(self.theta.device)
        x = torch.as_tensor(x, dtype=torch.float32, device=self.theta.device)
        single = x.dim() == 1
        pts = x.reshape(-1, 2).contiguous()
        u = K.field_eval(self.net_desc(K), self.theta.detach().contiguous(), pts, 1)[0]   # [N, out]
        if self.spec.sum_output:
            u = u.sum(dim=-1)
        return u[0] if single else u


def is_siren_model(obj):
    return isinstance(obj, Model)
