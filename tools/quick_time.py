"""Quick CUDA-event timing of the three jet kernels at C1- and C5-like shapes (development aid)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from meta_pde_b200 import ops, synthetic  # noqa: E402


def timeit(fn, warm=2, iters=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = torch.device("cuda:0")
    K = ops.kernels(dev)
    cases = [("poisson", 64, 3, 1024, 1), ("poisson", 64, 3, 1024, 16), ("poisson", 64, 3, 16384, 64),
             ("td_burgers", 64, 8, 1024, 8), ("hyper_elasticity", 64, 8, 1024, 8),
             ("poisson", 128, 5, 4096, 16), ("poisson", 256, 5, 16384, 8), ("poisson", 256, 5, 16384, 64)]
    if len(sys.argv) > 1:
        cases = [c for c in cases if str(c[1]) in sys.argv[1].split(",")]
    for pde, W, L, n, T in cases:
        out = synthetic.out_dim_of(pde)
        theta = synthetic.siren_theta([W] * L, out, 30.0, 30.0, True, seed=0)
        theta = theta.repeat(T, 1).contiguous().to(dev)
        sizes = synthetic.default_set_sizes(pde, n)
        sets, prm = synthetic.make_tasks(pde, T, sizes, seed=1)
        sets = [s.to(dev) for s in sets]
        prm = prm.to(dev)
        v = torch.randn_like(theta)
        net = K.net_desc(L, W, out, 30.0, True)
        Pw = 2 * W + (L - 1) * W * W + W * out
        C = 3 if pde == "hyper_elasticity" else 4
        R = C * sizes[-1] + sum(sizes[:-1])
        unit = 2.0 * Pw * R * T
        row = {"pde": pde, "W": W, "L": L, "n": n, "T": T}
        for name, mult, fn in [
            ("loss", 1, lambda: K.loss(net, pde, 1.0, theta, sets, prm, T)),
            ("grad", 3, lambda: K.loss_grad(net, pde, 1.0, theta, sets, prm, T)),
            ("hvp", 9, lambda: K.loss_hvp(net, pde, 1.0, theta, v, sets, prm, T)),
            ("hvp_nograd", 8, lambda: K.loss_hvp(net, pde, 1.0, theta, v, sets, prm, T, want_grad=False)),
        ]:
            ms = timeit(fn)
            row[name + "_ms"] = round(ms, 4)
            row[name + "_tflops"] = round(mult * unit / ms * 1e-9, 2)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
