"""the `--nosiren` network (nonlinearity = flax.nn.swish, no omega: src/maml_pde.py:115,
src/nets/field.py:187-188) through the same C ABI, kernels and checkers as the SIREN network.

CPU half: the kernel sources under the SIMT emulator (tests/emu) vs the float64 oracle.  GPU half (`-m gpu`): the CUDA
library; swish runs the FP32-FMA kernels at every width (the tcgen05 epilogues are sin-jet only), so widths 128 / 256 are
covered here on that path.  The reference-source swish fixtures (tests/golden/ref_*_swish.npz) run in test_golden.py."""
import ctypes as C

import pytest
import torch

import parity_cases as pc


# ------------------------------------------------------------------------------------------- emulator (CPU)
@pytest.mark.parametrize("pde,W,L,n", [
    ("poisson", 32, 2, 70),
    ("td_burgers", 32, 2, 70),
    ("hyper_elasticity", 32, 2, 50),
    ("poisson", 64, 3, 40),
])
def test_emu_loss_grad_hvp(emu_kernels, pde, W, L, n):
    with pc.nonlinearity("swish"):
        pc.check_loss_grad_hvp(emu_kernels, pde, W, L, n, T=2)


def test_emu_field_eval_and_rollouts(emu_kernels):
    with pc.nonlinearity("swish"):
        pc.check_field_eval(emu_kernels, 32, 2, 2, 130, T=1)
        pc.check_maml(emu_kernels, "poisson", 32, 2, 40, T=2, K_steps=2, inner_lr=1e-2, clip=1e-2, decay=0.5)
        pc.check_leap(emu_kernels, "td_burgers", 32, 2, 70, T=2, K_steps=2, inner_lr=1e-3)


def test_unknown_nonlinearity_is_refused(emu_kernels):
    K = emu_kernels
    net = K.net_desc(2, 32, 1, 30.0, True)
    net.nonlinearity = 7
    pde = K.pde_desc("poisson", 1.0, [8, 8])
    assert K.abi.mpde_workspace_bytes(C.byref(net), C.byref(pde), 1, 0) == 0     # size query: 0 = unsupported
    with pytest.raises(Exception):
        K.net_desc(2, 32, 1, 30.0, True, nonlinearity="relu")


def test_compat_field_accepts_swish():
    from meta_pde_b200.compat import field
    from meta_pde_b200.compat import random as jrandom
    spec = field.NeuralPotential.partial(sizes=[64, 64], nonlinearity=field.swish, omega=1.0, omega0=3.0, log_scale=True).spec()
    assert spec.nonlinearity_name == "swish"
    w = field.variance_scaling_fan_in_truncated_normal()(jrandom.PRNGKey(3), (64, 64))
    lim = 2.0 / 8.0 / 0.87962566103423978           # 2 sigma of N(0, 1/64), rescaled (field.py:171-176)
    assert w.shape == (64, 64) and abs(w).max() <= lim and 0.09 < w.std() < 0.16
    with pytest.raises(NotImplementedError):
        field.NeuralPotential.partial(sizes=[64], nonlinearity=field.swish, n_fourier=4).spec()


# ------------------------------------------------------------------------------------------- CUDA library
@pytest.mark.gpu
@pytest.mark.parametrize("pde,W,L,n", [
    ("poisson", 64, 3, 1024),            # the README command's shape with --nosiren
    ("td_burgers", 64, 3, 1024),
    ("hyper_elasticity", 64, 3, 1024),
    ("poisson", 32, 2, 70),
    ("poisson", 128, 2, 200),            # FP32-FMA kernels at the widths the SIREN network runs on tcgen05
    ("td_burgers", 256, 2, 100),
])
def test_gpu_loss_grad_hvp(gpu_kernels, pde, W, L, n):
    with pc.nonlinearity("swish"):
        pc.check_loss_grad_hvp(gpu_kernels, pde, W, L, n, T=2)


@pytest.mark.gpu
def test_gpu_field_eval_and_rollouts(gpu_kernels):
    with pc.nonlinearity("swish"):
        pc.check_field_eval(gpu_kernels, 64, 3, 1, 1024, T=3)
        pc.check_field_eval(gpu_kernels, 128, 2, 2, 500, T=1)
        pc.check_maml(gpu_kernels, "poisson", 64, 3, 128, T=3, K_steps=5, inner_lr=1e-2, clip=1e-2, decay=0.5)
        pc.check_maml(gpu_kernels, "hyper_elasticity", 64, 3, 64, T=2, K_steps=2, inner_lr=1e-4)
        pc.check_leap(gpu_kernels, "td_burgers", 64, 3, 256, T=2, K_steps=3, inner_lr=1e-3)
