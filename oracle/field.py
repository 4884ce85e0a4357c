"""Oracle: SIREN neural field (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py).

Follows /root/reference/src/nets/field.py:
  * siren_init / first_layer_siren_init            field.py:23-44
  * constant_init for the log io-scales            field.py:140-143
  * nf_apply                                       field.py:146-199
  * NeuralField1d (= NeuralPotential), make_nf_ndim field.py:212-226,259
  * laplace_operator / vmap_laplace_operator       field.py:47-72

Parameters live in a nested dict with the reference's (pre-Linen Flax) names:
``{'Dense_0': {'kernel': [n_in, n_out], 'bias': [n_out]}, ...,
   'log_input_scale': [1, in_dim], 'log_output_scale': [1, out_dim]}``.
"""
import math

import numpy as np
import torch
from torch.func import jacfwd, jacrev, vmap


# ----------------------------------------------------------------------------
# initialisers (field.py:23-44,140-143).  JAX's threefry stream is not
# reproduced here; the distributions are.
# ----------------------------------------------------------------------------
def siren_init(rng, shape, omega, dtype=np.float64):
    fan_in = shape[0]
    lim = math.sqrt(6.0 / fan_in) / omega
    return rng.uniform(-lim, lim, size=shape).astype(dtype)


def first_layer_siren_init(rng, shape, omega, omega0, dtype=np.float64):
    fan_in = shape[0]
    return ((omega0 / omega) * rng.uniform(-1.0 / fan_in, 1.0 / fan_in, size=shape)).astype(dtype)


def init_params(seed, sizes, in_dim=2, out_dim=1, omega=30.0, omega0=30.0,
                log_scale=True, io_scale_lr_factor=10.0, dtype=torch.float64,
                bias_scale=0.0):
    """Parameter tree in the creation order of nf_apply (field.py:160-197).

    ``bias_scale`` > 0 perturbs the (reference: zero) biases so that tests
    exercise the bias paths with non-trivial values.
    """
    rng = np.random.default_rng(seed)
    params = {}
    if log_scale:
        params["log_input_scale"] = torch.full((1, in_dim), math.log(1.0 / io_scale_lr_factor), dtype=dtype)
    n_in = in_dim
    for i, size in enumerate(sizes):
        if i == 0:
            k = first_layer_siren_init(rng, (n_in, size), omega, omega0)
        else:
            k = siren_init(rng, (n_in, size), omega)
        b = bias_scale * rng.standard_normal(size)
        params[f"Dense_{i}"] = {"kernel": torch.tensor(k, dtype=dtype), "bias": torch.tensor(b, dtype=dtype)}
        n_in = size
    k = siren_init(rng, (n_in, out_dim), omega)
    b = bias_scale * rng.standard_normal(out_dim)
    params[f"Dense_{len(sizes)}"] = {"kernel": torch.tensor(k, dtype=dtype), "bias": torch.tensor(b, dtype=dtype)}
    if log_scale:
        params["log_output_scale"] = torch.full((1, out_dim), math.log(1.0 / io_scale_lr_factor), dtype=dtype)
    return params


# ----------------------------------------------------------------------------
# flat theta layout of the C ABI (include/metapde_b200.h):
#   for l = 0..L: kernel_l [n_in, n_out] row-major, bias_l [n_out];
#   then log_input_scale [in_dim], log_output_scale [out_dim] (if log_scale)
# ----------------------------------------------------------------------------
def num_dense(params):
    return sum(1 for k in params if k.startswith("Dense_"))


def flatten_params(params):
    parts = []
    for l in range(num_dense(params)):
        parts.append(params[f"Dense_{l}"]["kernel"].reshape(-1))
        parts.append(params[f"Dense_{l}"]["bias"].reshape(-1))
    if "log_input_scale" in params:
        parts.append(params["log_input_scale"].reshape(-1))
        parts.append(params["log_output_scale"].reshape(-1))
    return torch.cat(parts)


def unflatten_params(theta, sizes, in_dim=2, out_dim=1, log_scale=True):
    params = {}
    off = 0
    n_in = in_dim
    dims = list(sizes) + [out_dim]
    if log_scale:
        # keep the reference creation order: log_input_scale first
        params["log_input_scale"] = None
    for l, n_out in enumerate(dims):
        k = theta[off:off + n_in * n_out].reshape(n_in, n_out)
        off += n_in * n_out
        b = theta[off:off + n_out]
        off += n_out
        params[f"Dense_{l}"] = {"kernel": k, "bias": b}
        n_in = n_out
    if log_scale:
        params["log_input_scale"] = theta[off:off + in_dim].reshape(1, in_dim)
        off += in_dim
        params["log_output_scale"] = theta[off:off + out_dim].reshape(1, out_dim)
        off += out_dim
    assert off == theta.numel(), (off, theta.numel())
    return params


def param_count(sizes, in_dim=2, out_dim=1, log_scale=True):
    n_in, p = in_dim, 0
    for n_out in list(sizes) + [out_dim]:
        p += n_in * n_out + n_out
        n_in = n_out
    if log_scale:
        p += in_dim + out_dim
    return p


# ----------------------------------------------------------------------------
# nf_apply (field.py:146-199): nonlinearity = np.sin (FLAGS.siren, the published runs; field.py:185-186 applies omega) or
# flax.nn.swish (`--nosiren`, maml_pde.py:115; field.py:187-188: no omega).  n_fourier / LAAF are never passed by a driver.
# ----------------------------------------------------------------------------
def nf_apply(params, out_dim, x, omega=30.0, nonlinearity="sin"):
    assert nonlinearity in ("sin", "swish"), nonlinearity
    log_scale = "log_input_scale" in params
    if log_scale:
        in_scale = torch.exp(params["log_input_scale"])
        x = (x.reshape(-1, x.shape[-1]) * in_scale).reshape(x.shape)
    n = num_dense(params)
    for i in range(n - 1):
        d = params[f"Dense_{i}"]
        a = x @ d["kernel"] + d["bias"]
        x = torch.sin(a * omega) if nonlinearity == "sin" else a * torch.sigmoid(a)
    d = params[f"Dense_{n - 1}"]
    out = x @ d["kernel"] + d["bias"]
    if log_scale:
        out_scale = torch.exp(params["log_output_scale"])
        out = (out.reshape(-1, out_dim) * out_scale).reshape(out.shape)
    return out


def neural_field_1d(params, x, omega=30.0, nonlinearity="sin"):
    """NeuralField1d / NeuralPotential (field.py:212-216,259)."""
    return nf_apply(params, 1, x, omega, nonlinearity).sum(dim=-1)


def neural_field_nd(params, x, n_dims, omega=30.0, nonlinearity="sin"):
    """make_nf_ndim(n_dims) (field.py:219-226)."""
    return nf_apply(params, n_dims, x, omega, nonlinearity)


def make_field_fn(params, out_dim, omega, nonlinearity="sin"):
    """The callable the reference hands to pde.loss_fn (a flax.nn.Model)."""
    if out_dim == 1:
        return lambda x: neural_field_1d(params, x, omega, nonlinearity)
    return lambda x: neural_field_nd(params, x, out_dim, omega, nonlinearity)


# ----------------------------------------------------------------------------
# laplace_operator (field.py:47-72): trace(jacfwd(x2 -> jacrev(u)(x2) * w(x2)))
# ----------------------------------------------------------------------------
def laplace_operator(x, potential_fn, weighting_fn=lambda x: 1.0):
    assert x.dim() == 1
    hess_fn = jacfwd(lambda x2: jacrev(lambda x1: potential_fn(x1).squeeze())(x2) * weighting_fn(x2))
    hess = hess_fn(x)
    assert hess.dim() == 2
    return torch.trace(hess)


def vmap_laplace_operator(x, potential_fn, weighting_fn=lambda x: 1.0):
    return vmap(lambda xi: laplace_operator(xi, potential_fn, weighting_fn))(x)
