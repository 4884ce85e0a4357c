#!/bin/bash
# round 2, pass g: ncu --set full of the --nosiren (swish) FP32-FMA kernels at the published 3x64 network, saturating grid
# (64 tasks x 16384 + 16384 points), GRAD and HVP modes; CUDA-event timing of sin vs swish at the same shape beside it.
export MPDE_W=64 MPDE_L=3 MPDE_T=64 MPDE_N=16384 MPDE_ACT=swish
bash tools/gpu_ncu_full.sh r02_g_swish grad pde_kernel_swish hvp pde_kernel_swish
python - <<'P' > gpurun_out/r02_g_swish_vs_sin_time.jsonl 2>&1
import json, os, sys, torch
sys.path.insert(0, os.getcwd())
from meta_pde_b200 import ops, synthetic
K = ops.kernels(torch.device("cuda:0")); dev = K.device
W, L, T, n = 64, 3, 64, 16384
sets, prm = synthetic.make_tasks("poisson", T, [n, n], seed=1); sets = [s.to(dev) for s in sets]; prm = prm.to(dev)
for act in ("sin", "swish"):
    th = synthetic.siren_theta([W] * L, 1, 30.0, 30.0, True, seed=0) if act == "sin" else synthetic.swish_theta([W] * L, 1, True, seed=0)
    th = th.to(dev).unsqueeze(0).repeat(T, 1).contiguous(); v = torch.randn_like(th)
    net = K.net_desc(L, W, 1, 30.0, True, nonlinearity=act)
    fns = {"loss": lambda: K.loss(net, "poisson", 1.0, th, sets, prm, T), "grad": lambda: K.loss_grad(net, "poisson", 1.0, th, sets, prm, T),
           "hvp": lambda: K.loss_hvp(net, "poisson", 1.0, th, v, sets, prm, T)}
    for mode, fn in fns.items():
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): fn()
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"act": act, "mode": mode, "W": W, "L": L, "T": T, "n": n, "ms_per_call": e0.elapsed_time(e1) / 10}))
P
cat gpurun_out/r02_g_swish_vs_sin_time.jsonl
