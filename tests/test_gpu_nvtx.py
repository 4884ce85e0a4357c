"""MPDE_NVTX=1 (NVTX ranges around the ABI entry points, api.cu) must not change results: the same seeded loss+grad call
in a child process with the ranges on and in this process with them off (the flag is read once per process)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, torch
sys.path.insert(0, %r)
from meta_pde_b200 import ops, synthetic
K = ops.kernels(torch.device("cuda", 0))
theta = synthetic.siren_theta([64] * 3, 1, 30.0, 30.0, True, seed=3).to(K.device)
sets, prm = synthetic.make_tasks("poisson", 2, [100, 100], seed=4)
net = K.net_desc(3, 64, 1, 30.0, True)
loss, terms, grad = K.loss_grad(net, "poisson", 1.0, theta, [s.to(K.device) for s in sets], prm.to(K.device), 2)
u = K.field_eval(net, theta.unsqueeze(0).contiguous(), sets[1][:1].to(K.device), 1)
torch.cuda.synchronize()
print("RESULT", float(loss.double().sum()), float(grad.double().abs().sum()), float(u.double().sum()))
"""


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    out = subprocess.run([sys.executable, "-c", CHILD % ROOT], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("RESULT")][-1]
    return [float(x) for x in line.split()[1:]]


def test_nvtx_ranges_do_not_change_results(gpu_kernels):
    assert _run({"MPDE_NVTX": "1"}) == _run({"MPDE_NVTX": "0"})
