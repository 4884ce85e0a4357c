"""One call of a single mode at C5 size (for ncu captures): python tools/ncu_one.py grad|hvp|loss [reps]
(MPDE_W / MPDE_L / MPDE_T / MPDE_N override the 5x256, 64 tasks, 16384 points; MPDE_ACT=swish: the --nosiren network)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from meta_pde_b200 import ops, synthetic

mode = sys.argv[1] if len(sys.argv) > 1 else "grad"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
K = ops.kernels(torch.device("cuda:0"))
dev = K.device
W, L = int(os.environ.get("MPDE_W", "256")), int(os.environ.get("MPDE_L", "5"))
T, n = int(os.environ.get("MPDE_T", "64")), int(os.environ.get("MPDE_N", "16384"))
ACT = os.environ.get("MPDE_ACT", "sin")
theta0 = synthetic.siren_theta([W] * L, 1, 30.0, 30.0, True, seed=0) if ACT == "sin" else synthetic.swish_theta([W] * L, 1, True, seed=0)
theta = theta0.to(dev).unsqueeze(0).repeat(T, 1).contiguous()
sets, prm = synthetic.make_tasks("poisson", T, [n, n], seed=1)
sets = [s.to(dev) for s in sets]
prm = prm.to(dev)
net = K.net_desc(L, W, 1, 30.0, True, nonlinearity=ACT)
v = torch.randn_like(theta)
fn = {"loss": lambda: K.loss(net, "poisson", 1.0, theta, sets, prm, T),
      "grad": lambda: K.loss_grad(net, "poisson", 1.0, theta, sets, prm, T),
      "hvp": lambda: K.loss_hvp(net, "poisson", 1.0, theta, v, sets, prm, T)}[mode]
for _ in range(reps):
    out = fn()
torch.cuda.synchronize()
print(mode, float(out[0][0]))
