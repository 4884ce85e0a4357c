"""Block-wise comparison of the tensor-core gradient (loss_grad) with the FP32-FMA gradient (from loss_hvp)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import parity_cases as pc
from meta_pde_b200 import ops
from meta_pde_b200._abi import PDE_OUT_DIM

K = ops.kernels(torch.device("cuda:0"))
for pde, W, L, n, T in [("poisson", 256, 3, 300, 1), ("poisson", 256, 3, 3000, 1), ("poisson", 256, 3, 16384, 1), ("hyper_elasticity", 128, 3, 100, 1)]:
    theta, sets, prm = pc.make_problem(pde, W, L, n, T, 0, True, None)
    OUT = PDE_OUT_DIM[pde]
    net = K.net_desc(L, W, OUT, 30.0, True)
    d = lambda x: x.to(K.device)
    dsets = [d(s) for s in sets]
    v = torch.randn_like(theta)
    _, _, g1 = K.loss_grad(net, pde, 1.0, d(theta), dsets, d(prm), T)
    g1 = g1.clone().cpu()
    _, _, g2, _ = K.loss_hvp(net, pde, 1.0, d(theta), d(v), dsets, d(prm), T)
    g2 = g2.clone().cpu()
    blocks = [("w0", 0, 2 * W), ("b0", 2 * W, 3 * W)]
    off = 3 * W
    for l in range(1, L):
        blocks.append((f"W{l}", off, off + W * W)); off += W * W
        blocks.append((f"b{l}", off, off + W)); off += W
    blocks.append(("wL", off, off + W * OUT)); off += W * OUT
    blocks.append(("bL", off, off + OUT)); off += OUT
    blocks.append(("log", off, g1.shape[1]))
    for t in range(T):
        out = {}
        for name, a, b in blocks:
            x, y = g1[t, a:b].double(), g2[t, a:b].double()
            out[name] = [float(f"{float((x - y).norm() / y.norm().clamp_min(1e-30)):.2e}"), float(f"{float(x.norm() / y.norm().clamp_min(1e-30)):.3f}")]
        print(json.dumps({"case": [pde, W, L, n], "task": t, "rel_err,norm_ratio": out}), flush=True)
