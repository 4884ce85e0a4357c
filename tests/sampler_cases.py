"""Shared checks of the GPU samplers (csrc/sampler_kernels.cuh + compat/gpu_samplers.py) against the host samplers of
compat/pdes.py -- which tests/test_sampler_parity.py pins to the reference's own sampler source.  Used with the emulator
build on the CPU (tests/test_emu_kernels.py) and with the CUDA library on the GPU (tests/test_gpu_compat.py)."""
import numpy as np
import torch

from meta_pde_b200.compat import flags as cflags

cflags.define_maml()
cflags.ensure_parsed(["test"])
from meta_pde_b200.compat import gpu_samplers as gs, pdes, random as jr  # noqa: E402


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def check_threefry(K):
    keys = np.stack([jr.PRNGKey(0), jr.PRNGKey(7), jr.split(jr.PRNGKey(3))[1], np.array([0xFFFFFFFF, 0x12345678], np.uint32)])
    # the reference's own jax.random known answer (src/.ipynb_checkpoints/Untitled-checkpoint.ipynb:183)
    sp = _np(K.random_split(keys[:1], 3)).view(np.uint32)[0]
    assert sp.tolist() == [[2467461003, 428148500], [3186719485, 3840466878], [2562233961, 1946702221]]
    for n in (1, 2, 5, 8, 4097):                         # odd sizes: the zero-padded counter path
        bits = _np(K.random_bits(keys, n)).view(np.uint32)
        assert (bits == np.stack([jr.random_bits(k, (n,)) for k in keys])).all(), n
        u = _np(K.random_uniform(keys, n, -0.3, 2.5))
        assert (u == np.stack([jr.uniform(k, (n,), minval=-0.3, maxval=2.5) for k in keys])).all(), n   # bit-exact
    for num in (2, 3, 5, 7):
        sp = _np(K.random_split(keys, num)).view(np.uint32)
        assert (sp == np.stack([jr.split(k, num) for k in keys])).all(), num


def check_poisson(K, n=64, S=6):
    p = pdes.get_pde("poisson")
    keys = jr.split(jr.PRNGKey(5), S)
    prms = [p.sample_params(k) for k in jr.split(jr.PRNGKey(9), S)]
    b, d = gs.sample_points_batched(p, keys, n, prms, K)
    assert tuple(b.shape) == (S, n, 2) and tuple(d.shape) == (S, n, 2)
    rows_equal = 0
    for s in range(S):
        hb, hd = (_np(x) for x in p.sample_points(keys[s], n, prms[s]))
        np.testing.assert_allclose(_np(b[s]), hb, rtol=0, atol=2e-6)                  # cos / sin: ulp-level differences
        rows_equal += int((np.abs(_np(d[s]) - hd).max(axis=1) < 1e-6).sum())
        # every selected point is admissible (outside the hole), whatever the tie-breaking
        c1, c2 = (float(v) for v in _np(prms[s][2]))
        x, y = _np(d[s])[:, 0], _np(d[s])[:, 1]
        th = np.arctan2(x, y)
        assert ((1.0 + c1 * np.cos(4 * th) + c2 * np.cos(8 * th)) >= np.sqrt(x * x + y * y) - 1e-5).all()
    # same candidates in the same order, except where two Gumbel keys differ by an ulp (none expected at this size)
    assert rows_equal >= int(0.98 * S * n), (rows_equal, S * n)


def check_burgers(K, n=200, S=5):
    q = pdes.get_pde("td_burgers")
    keys = jr.split(jr.PRNGKey(11), S)
    sets = gs.sample_points_batched(q, keys, n, [None] * S, K)
    prm = q.sample_params(jr.PRNGKey(1))
    for s in range(S):
        host = q.sample_points(keys[s], n, prm)
        for a, h in zip(sets, host):
            assert np.array_equal(_np(a[s]), _np(h))                                   # uniform draws only: bit-exact
