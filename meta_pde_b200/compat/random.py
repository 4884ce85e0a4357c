"""``jax.random`` restated for the host side of the plugin surface (numpy, uint32 arithmetic).

The reference draws task parameters and collocation points with ``jax.random`` (jax 0.2.20, threefry2x32;
``src/poisson/poisson_common.py:42-119``, ``src/burgers/td_burgers_common.py:94-186``,
``src/elasticity/hyper_elasticity_common.py:89-363``).  jax is a third-party dependency absent from
/root/reference and from this image; this module restates the published Threefry-2x32 counter PRNG
(Salmon et al., SC'11; 20 rounds, rotations 13/15/26/6, 17/29/16/24, key-schedule constant 0x1BD11BDA) and the
jax 0.2.20 derivations on top of it (``split``: counts iota(2n); ``_random_bits``: counts iota(size), odd sizes
zero-padded, the counter array split in two halves; ``uniform``: mantissa trick on bits >> 9).

Pinned by the only golden vectors the reference holds for anything near the path
(``src/.ipynb_checkpoints/Untitled-checkpoint.ipynb:183,215,236,257-259``): see tests/test_compat_host.py::test_prng_golden_vectors (and tests/test_sampler_parity.py for the samplers built on it).
Sampling is host-side plumbing: the kernels consume its output.
"""
import numpy as np

_U32 = np.uint32


def _rotl(x, r):
    return (x << _U32(r)) | (x >> _U32(32 - r))


def threefry2x32(key, x0, x1):
    """One Threefry-2x32-20 block per element of the uint32 arrays (x0, x1) under the key pair."""
    k0, k1 = _U32(key[0]), _U32(key[1])
    ks = [k0, k1, k0 ^ k1 ^ _U32(0x1BD11BDA)]
    rot = [(13, 15, 26, 6), (17, 29, 16, 24)]
    with np.errstate(over="ignore"):
        x0 = x0.astype(_U32) + ks[0]
        x1 = x1.astype(_U32) + ks[1]
        for i in range(5):
            for r in rot[i % 2]:
                x0 = x0 + x1
                x1 = _rotl(x1, r)
                x1 = x1 ^ x0
            x0 = x0 + ks[(i + 1) % 3]
            x1 = x1 + ks[(i + 2) % 3] + _U32(i + 1)
    return x0, x1


def _threefry_counts(key, counts):
    counts = np.asarray(counts, dtype=_U32).ravel()
    n = counts.size
    if n % 2:
        counts = np.concatenate([counts, np.zeros(1, _U32)])
    h = counts.size // 2
    a, b = threefry2x32(key, counts[:h], counts[h:])
    return np.concatenate([a, b])[:n]


def PRNGKey(seed):
    seed = int(seed)
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=_U32)


def split(key, num=2):
    return _threefry_counts(key, np.arange(2 * num, dtype=_U32)).reshape(num, 2)


def random_bits(key, shape):
    size = int(np.prod(shape)) if len(shape) else 1
    return _threefry_counts(key, np.arange(size, dtype=_U32)).reshape(shape)


def uniform(key, shape=(), dtype=np.float32, minval=0.0, maxval=1.0):
    shape = tuple(shape)
    bits = random_bits(key, shape)
    floats = ((bits >> _U32(9)) | _U32(0x3F800000)).view(np.float32) - np.float32(1.0)
    lo, hi = np.float32(minval), np.float32(maxval)
    return np.maximum(lo, floats * (hi - lo) + lo).astype(np.float32)


def normal(key, shape=(), dtype=np.float32):
    from scipy.special import erfinv

    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    u = uniform(key, shape, minval=lo, maxval=1.0)
    return (np.float32(np.sqrt(2.0)) * erfinv(u.astype(np.float64))).astype(np.float32)


def gumbel(key, shape=()):
    u = uniform(key, shape, minval=np.finfo(np.float32).tiny, maxval=1.0)
    return -np.log(-np.log(u))


def randint(key, shape, minval, maxval):
    """jax 0.2.20 randint for spans far below 2**32: two 32-bit draws combined modulo the span."""
    k1, k2 = split(key)
    hi_bits = random_bits(k1, tuple(shape)).astype(np.uint64)
    lo_bits = random_bits(k2, tuple(shape)).astype(np.uint64)
    span = np.uint64(int(maxval) - int(minval))
    mult = (np.uint64(2 ** 16) % span) ** np.uint64(2) % span
    off = ((hi_bits % span) * mult + (lo_bits % span)) % span
    return (int(minval) + off.astype(np.int64)).astype(np.int32)


def choice_without_replacement(key, n_inputs, n_draws, p):
    """jax.random.choice(key, n, shape=(n_draws,), replace=False, p=p): Gumbel top-k (argsort of -g - log p)."""
    p = np.asarray(p, dtype=np.float32)
    with np.errstate(divide="ignore"):
        g = -gumbel(key, (n_inputs,)) - np.log(p)
    return np.argsort(g, kind="stable")[:n_draws]
