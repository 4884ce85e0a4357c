"""Point sampling for one outer step, host samplers vs GPU samplers (compat/gpu_samplers.py): the (2K + 2) x T point sets of
C5 (Poisson, 64 tasks, K = 5, 16384 points per set) and of the published 3x64 shape (16 tasks x 1024).  One JSON line each."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from meta_pde_b200.compat import flags as cflags  # noqa: E402

cflags.define_maml()
cflags.ensure_parsed(["sampler_time"])
from meta_pde_b200 import ops  # noqa: E402
from meta_pde_b200.compat import gpu_samplers as gs, pdes, random as jr  # noqa: E402

K = ops.kernels(torch.device("cuda:0"))
for name, T, Ksteps, n in (("poisson", 64, 5, 16384), ("poisson", 16, 5, 1024), ("td_burgers", 8, 5, 1024)):
    pde = pdes.get_pde(name)
    S = (2 * Ksteps + 2) * T
    keys = jr.split(jr.PRNGKey(0), S)
    prms = [pde.sample_params(k) for k in jr.split(jr.PRNGKey(1), T)] * (2 * Ksteps + 2)
    gs.sample_points_batched(pde, keys[:4], n, prms[:4], K)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sets = gs.sample_points_batched(pde, keys, n, prms, K)
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0
    m = min(S, 8)
    t0 = time.perf_counter()
    for s in range(m):
        pde.sample_points(keys[s], n, prms[s])
    t_host = (time.perf_counter() - t0) * S / m
    print(json.dumps({"what": "sampling one outer step", "pde": name, "tasks": T, "inner_steps": Ksteps, "points_per_set": n,
                      "point_sets": S, "gpu_samplers_s": t_gpu, "host_samplers_s": t_host,
                      "host_note": f"{m} of {S} samples timed, scaled", "bytes_sampled": int(sum(x.numel() * 4 for x in sets))}), flush=True)
