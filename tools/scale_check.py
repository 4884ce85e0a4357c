"""Full-size robustness check of the tensor-core path on the other two PDEs: loss / grad / Hv at T tasks x n points,
finite outputs, H v linear in v, grad of the HVP call equal to the GRAD call.  usage: python tools/scale_check.py [T] [n]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from meta_pde_b200 import ops, synthetic

T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
K = ops.kernels(torch.device("cuda:0"))
dev = K.device
for pde, W, L, out in (("td_burgers", 256, 5, 1), ("hyper_elasticity", 256, 5, 2), ("td_burgers", 128, 8, 1)):
    theta = synthetic.siren_theta([W] * L, out, 30.0, 30.0, True, seed=0).to(dev).unsqueeze(0).repeat(T, 1).contiguous()
    sizes = synthetic.default_set_sizes(pde, n)
    sets, prm = synthetic.make_tasks(pde, T, sizes, seed=1)
    sets = [s.to(dev) for s in sets]
    prm = prm.to(dev)
    net = K.net_desc(L, W, out, 30.0, True)
    g = torch.Generator(device="cpu").manual_seed(3)
    v1 = torch.randn(theta.shape, generator=g).to(dev)
    v2 = torch.randn(theta.shape, generator=g).to(dev)
    l0, _, g0 = K.loss_grad(net, pde, 1.0, theta, sets, prm, T)
    l1, _, g1, h1 = K.loss_hvp(net, pde, 1.0, theta, v1, sets, prm, T)
    _, _, _, h2 = K.loss_hvp(net, pde, 1.0, theta, v2, sets, prm, T)
    _, _, _, h12 = K.loss_hvp(net, pde, 1.0, theta, (v1 + 0.5 * v2).contiguous(), sets, prm, T)
    torch.cuda.synchronize()
    rel = lambda a, b: float((a - b).norm() / b.norm())
    sym = float(((h1 * v2).sum(1) - (h2 * v1).sum(1)).abs().max() / (h1 * v2).sum(1).abs().max())
    rec = {"pde": pde, "net": f"{L}x{W}", "T": T, "sizes": sizes, "finite": bool(torch.isfinite(g0).all() and torch.isfinite(h1).all()),
           "loss_grad_vs_hvp_call": rel(g1, g0), "loss_match": rel(l1, l0), "hv_linearity": rel(h12, h1 + 0.5 * h2),
           "hv_symmetry": sym}
    print(json.dumps(rec), flush=True)
    assert rec["finite"] and rec["loss_grad_vs_hvp_call"] < 1e-4 and rec["hv_linearity"] < 1e-4 and rec["hv_symmetry"] < 1e-3, rec
print("scale check ok")
