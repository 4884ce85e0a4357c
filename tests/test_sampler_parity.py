"""compat samplers / initialisers vs the REFERENCE'S OWN sampler and initialiser source.

tests/golden/ref_samplers.npz and ref_inits.npz were produced by tests/golden/make_sampler_fixtures.py, which executes
/root/reference/src/{poisson/poisson_common,burgers/td_burgers_common,elasticity/hyper_elasticity_common}.py
(sample_params, sample_points, sample_points_in_domain) and src/nets/field.py (siren_init, first_layer_siren_init,
constant_init) under the jax stand-in with the Threefry PRNG of compat/random.py plugged in as jax.random.
Keys and drawn indices must agree exactly (the same rows are selected); floats to 1e-6 (the stand-in evaluates the
reference's float32 formulas in float64).  SURVEY.md section 8 rows a2, a8, f3.
"""
import ast
import os

import numpy as np
import pytest
import torch

from meta_pde_b200.compat import flags as cflags

cflags.define_maml()
cflags.ensure_parsed(["test"])
from meta_pde_b200.compat import field, pdes, random as jr  # noqa: E402
from meta_pde_b200.compat.flags import FLAGS  # noqa: E402

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SAMPLERS = np.load(os.path.join(G, "ref_samplers.npz"), allow_pickle=False)
INITS = np.load(os.path.join(G, "ref_inits.npz"), allow_pickle=False)
ATOL = 1e-6


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def _cases():
    for m in SAMPLERS["meta"]:
        tag, pde, seed, n, npar, npts, over = str(m).split(";", 6)
        yield pytest.param(tag, pde, int(seed), int(n), int(npar), int(npts), dict(ast.literal_eval(over)), id=f"{tag}-{pde}-n{n}")


@pytest.mark.parametrize("tag,pde,seed,n,npar,npts,over", list(_cases()))
def test_samplers_match_reference_source(tag, pde, seed, n, npar, npts, over):
    saved = {k: getattr(FLAGS, k) for k in over}
    for k, v in over.items():
        setattr(FLAGS, k, v)
    try:
        p = pdes.get_pde(pde)
        k0 = jr.PRNGKey(seed)
        assert k0.tolist() == SAMPLERS[f"{tag}_key"].tolist()
        kp, kx = jr.split(k0)
        params = p.sample_params(kp)
        for i in range(npar):
            ref = SAMPLERS[f"{tag}_param{i}"]
            got = _np(params[i]) if not isinstance(params[i], int) else np.asarray(params[i])
            assert got.reshape(-1).shape == ref.reshape(-1).shape, (i, got.shape, ref.shape)
            np.testing.assert_allclose(got.reshape(-1).astype(np.float64), ref.reshape(-1), rtol=0, atol=ATOL * max(1.0, np.abs(ref).max()),
                                       err_msg=f"{pde} param {i}")
        points = p.sample_points(kx, n, params)
        assert len(points) == npts
        for i in range(npts):
            ref = SAMPLERS[f"{tag}_points{i}"]
            got = _np(points[i]).astype(np.float64)
            assert got.shape == ref.shape, (i, got.shape, ref.shape)
            # the same candidate rows were selected (index parity) and their coordinates agree
            np.testing.assert_allclose(got, ref, rtol=0, atol=ATOL * max(1.0, np.abs(ref).max()), err_msg=f"{pde} point set {i}")
        dom = _np(p.sample_points_in_domain(kx, n, params)).astype(np.float64)
        np.testing.assert_allclose(dom, SAMPLERS[f"{tag}_domain"], rtol=0, atol=ATOL * 4, err_msg=f"{pde} sample_points_in_domain")
    finally:
        for k, v in saved.items():
            setattr(FLAGS, k, v)


@pytest.mark.parametrize("m", [str(x) for x in INITS["meta"]])
def test_siren_inits_match_reference_source(m):
    tag, seed, r, c, omega, omega0 = m.split(";")
    shape, omega, omega0 = (int(r), int(c)), float(omega), float(omega0)
    k = jr.PRNGKey(int(seed))
    a = field.siren_init(omega)(k, shape)
    b = field.first_layer_siren_init(omega, omega0)(k, shape)
    c_ = field.constant_init(-1.25)(k, (1, shape[1]))
    assert a.dtype == np.float32 and b.dtype == np.float32
    for got, name in ((a, "siren"), (b, "first"), (c_, "const")):
        ref = INITS[f"{tag}_{name}"]
        assert got.shape == ref.shape
        np.testing.assert_allclose(got.astype(np.float64), ref, rtol=2e-7, atol=1e-9, err_msg=name)
    lim = np.sqrt(6.0 / shape[0]) / omega
    assert np.abs(a).max() <= lim and np.abs(b).max() <= (omega0 / omega) / shape[0]


def test_init_by_shape_uses_the_pinned_initialisers():
    """init_by_shape = one threefry subkey per parameter in creation order through the three initialisers above."""
    cls = field.NeuralPotential.partial(sizes=[32, 32], dense_args=(), nonlinearity=field.sin, omega=30.0, omega0=30.0, log_scale=True)
    key = jr.PRNGKey(5)
    if torch.cuda.is_available():
        pytest.skip("host-only check (init_by_shape evaluates the field on the GPU when there is one)")
    try:
        _, params = cls.init_by_shape(key, [((1, 2), np.float32)])
    except Exception as e:   # no CUDA library on this host: Model() needs the kernels only when called
        pytest.skip(f"Model construction needs the CUDA build: {e}")
    shapes = cls.spec().param_shapes()
    keys = jr.split(key, len(shapes))
    k0 = _np(params["Dense_0"]["kernel"])
    np.testing.assert_array_equal(k0, field.first_layer_siren_init(30.0, 30.0)(keys[0], (2, 32)))
    k1 = _np(params["Dense_1"]["kernel"])
    np.testing.assert_array_equal(k1, field.siren_init(30.0)(keys[2], (32, 32)))
    assert float(_np(params["log_input_scale"]).reshape(-1)[0]) == pytest.approx(np.log(1.0 / FLAGS.io_scale_lr_factor), rel=1e-6)
