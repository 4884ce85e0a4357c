#!/usr/bin/env python
"""Generate tests/golden/ref_samplers.npz and ref_inits.npz by EXECUTING THE REFERENCE'S OWN sampler / initialiser
source (read from /root/reference, never copied) under the torch-backed jax stand-in of tests/golden/jaxshim.py, with
the stand-in's PRNG swapped for the Threefry-2x32 restatement of jax 0.2.20's jax.random
(meta_pde_b200/compat/random.py, itself pinned by the reference's three jax.random known answers).

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_sampler_fixtures.py

What runs from the reference, unmodified:
    src/nets/field.py:23-44,140-143            siren_init, first_layer_siren_init, constant_init
    src/poisson/poisson_common.py:42-119       sample_params, sample_points (boundary + domain incl. the arctan2 quirk)
    src/burgers/td_burgers_common.py:94-186    sample_params, sample_points (left/right share k2), sample_time
    src/elasticity/hyper_elasticity_common.py:89-363  sample_params (feasibility while_loop), sample_points (6 sets)
What this pins: the key-splitting order, which key feeds which draw, the rejection / masking logic and the geometry of
every sampler, i.e. everything the reference's own source decides.  What it cannot pin: jax.random itself (third-party,
absent) -- that is compat/random.py's restatement, shared by both sides of the comparison.

tests/test_sampler_parity.py compares meta_pde_b200/compat/pdes.py and compat/field.py with these arrays: keys and
indices bit-exact, floats to 1e-6 (the stand-in evaluates the reference's float32 formulas in float64).
"""
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
from meta_pde_b200.compat import random as jr  # noqa: E402  (numpy only)

jaxshim.use_threefry(jr)
sys.path.insert(0, REF)

from absl import flags  # noqa: E402
from src.util import common_flags  # noqa: E402,F401

FLAGS = flags.FLAGS
FLAGS(["fixtures"])

from src.nets import field as ref_field  # noqa: E402
from src.poisson import poisson_common as ref_poisson  # noqa: E402
from src.burgers import td_burgers_common as ref_burgers  # noqa: E402
from src.elasticity import hyper_elasticity_common as ref_hyper  # noqa: E402


def arr(x):
    if torch.is_tensor(x):
        return x.detach().cpu().numpy()
    return onp.asarray(x)


def key_of(seed):
    return jax.random.PRNGKey(seed)


def main():
    out = {}
    meta = []
    # ---- samplers: (pde, seed, n, flag overrides) ----
    cases = [("poisson", 0, 16, {}), ("poisson", 7, 64, {}), ("poisson", 3, 33, {"vary_geometry": False, "bc_scale": 2.0}),
             ("td_burgers", 0, 128, {}), ("td_burgers", 5, 200, {"vary_ic": False}),
             ("hyper_elasticity", 0, 16, {"max_holes": 2}), ("hyper_elasticity", 4, 24, {"max_holes": 3, "max_hole_size": 0.3}),
             ("poisson", 11, 16, {"fixed_num_pdes": 4, "seed": 9})]
    mods = {"poisson": ref_poisson, "td_burgers": ref_burgers, "hyper_elasticity": ref_hyper}
    for ci, (pde, seed, n, over) in enumerate(cases):
        saved = {k: getattr(FLAGS, k) for k in over}
        for k, v in over.items():
            setattr(FLAGS, k, v)
        try:
            mod = mods[pde]
            k0 = key_of(seed)
            kp, kx = jax.random.split(k0)
            params = mod.sample_params(kp)
            points = mod.sample_points(kx, n, params)
            dom = mod.sample_points_in_domain(kx, n, params)
            tag = f"c{ci}"
            out[f"{tag}_key"] = arr(k0).astype(onp.uint32)
            for i, p_ in enumerate(params):
                out[f"{tag}_param{i}"] = arr(p_).astype(onp.float64)
            for i, p_ in enumerate(points):
                out[f"{tag}_points{i}"] = arr(p_).astype(onp.float64)
            out[f"{tag}_domain"] = arr(dom).astype(onp.float64)
            meta.append((tag, pde, seed, n, len(params), len(points), repr(sorted(over.items()))))
            print(tag, pde, seed, n, over, [tuple(arr(p_).shape) for p_ in points])
        finally:
            for k, v in saved.items():
                setattr(FLAGS, k, v)
    out["meta"] = onp.array(meta, dtype=object)
    onp.savez_compressed(os.path.join(HERE, "ref_samplers.npz"), **{k: v for k, v in out.items() if k != "meta"},
                         meta=onp.array([";".join(str(x) for x in m) for m in meta]))

    # ---- initialisers ----
    ini = {}
    imeta = []
    for ii, (seed, shape, omega, omega0) in enumerate([(0, (2, 64), 30.0, 30.0), (1, (64, 64), 30.0, 30.0), (2, (256, 256), 30.0, 30.0),
                                                       (3, (2, 32), 10.0, 3.0), (4, (128, 1), 30.0, 30.0)]):
        k = key_of(seed)
        ini[f"i{ii}_siren"] = arr(ref_field.siren_init(omega)(k, shape)).astype(onp.float64)
        ini[f"i{ii}_first"] = arr(ref_field.first_layer_siren_init(omega, omega0)(k, shape)).astype(onp.float64)
        ini[f"i{ii}_const"] = arr(ref_field.constant_init(-1.25)(k, (1, shape[1]))).astype(onp.float64)
        imeta.append(f"i{ii};{seed};{shape[0]};{shape[1]};{omega};{omega0}")
    onp.savez_compressed(os.path.join(HERE, "ref_inits.npz"), **ini, meta=onp.array(imeta))
    print("wrote ref_samplers.npz, ref_inits.npz")


if __name__ == "__main__":
    main()
