"""Quick GPU check of the tensor-core path: parity of the routed modes vs the float64 oracle, then timing at C5 size.
usage: python tools/tc_check.py [parity|time] ...   (MPDE_DISABLE_TC=1 selects the FP32-FMA kernels for the same call)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

from meta_pde_b200 import ops, synthetic


def parity():
    import parity_cases as pc
    K = ops.kernels(torch.device("cuda:0"))
    for pde, W, L, n in [("poisson", 128, 2, 200), ("poisson", 256, 5, 300), ("td_burgers", 256, 2, 100),
                         ("hyper_elasticity", 128, 4, 150), ("poisson", 256, 3, 1000), ("td_burgers", 128, 8, 256)]:
        try:
            errs = pc.check_loss_grad_hvp(K, pde, W, L, n, T=2)
            print(json.dumps({"case": [pde, W, L, n], "ok": True, "errs": {k: float(f"{v:.3e}") for k, v in errs.items()}}), flush=True)
        except AssertionError as e:
            print(json.dumps({"case": [pde, W, L, n], "ok": False, "msg": str(e)[:1500]}), flush=True)


def timing(modes=("loss",)):
    K = ops.kernels(torch.device("cuda:0"))
    dev = K.device
    TT = int(os.environ.get("MPDE_T", "64"))
    for (W, L, T, n) in [(256, 5, TT, 16384), (128, 5, TT, 16384)]:
        theta = synthetic.siren_theta([W] * L, 1, 30.0, 30.0, True, seed=0).to(dev)
        theta = theta.unsqueeze(0).repeat(T, 1).contiguous()
        sets, prm = synthetic.make_tasks("poisson", T, [n, n], seed=1)
        sets = [s.to(dev) for s in sets]; prm = prm.to(dev)
        net = K.net_desc(L, W, 1, 30.0, True)
        v = torch.randn_like(theta)
        Pw = 2 * W + (L - 1) * W * W + W
        rows = 5 * n
        for mode in modes:
            fn = {"loss": lambda: K.loss(net, "poisson", 1.0, theta, sets, prm, T),
                  "grad": lambda: K.loss_grad(net, "poisson", 1.0, theta, sets, prm, T),
                  "hvp": lambda: K.loss_hvp(net, "poisson", 1.0, theta, v, sets, prm, T)}[mode]
            mult = {"loss": 2, "grad": 6, "hvp": 18}[mode]
            for _ in range(2):
                out = fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            K.prof_enable(True)
            e0.record()
            for _ in range(reps):
                out = fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            mid = {"loss": 0, "grad": 1, "hvp": 2}[mode]
            main_ms, main_n = K.prof_read(mid)
            wg_ms, wg_n = K.prof_read(3)
            K.prof_enable(False)
            print(json.dumps({"W": W, "L": L, "T": T, "n": n, "mode": mode, "tc": os.environ.get("MPDE_DISABLE_TC") != "1",
                              "ms": round(ms, 3), "alg_tflops": round(mult * Pw * rows * T / ms * 1e-9, 2),
                              "main_kernel_ms": round(main_ms / max(main_n, 1), 3), "wgrad_ms_per_launch": round(wg_ms / max(wg_n, 1), 3),
                              "wgrad_launches_per_call": wg_n / reps, "loss0": float(out[0][0])}), flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "parity"
    if what == "parity":
        parity()
    else:
        timing(tuple(sys.argv[2:]) or ("loss",))
