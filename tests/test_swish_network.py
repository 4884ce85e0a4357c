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


def _compat_swish_loss_vs_oracle(K, device):
    """The plugin surface with --nosiren: pde.BaseField.partial(nonlinearity=field.swish) -> Model -> pde.loss_fn (one fused
    launch, autograd to first order through mpde_loss_grad) and field_fn(x), against the float64 oracle."""
    from helpers import rel_l2
    from meta_pde_b200.compat import flags as cflags
    cflags.define_maml()             # idempotent: the other compat test modules do the same at import
    cflags.ensure_parsed(["test"])
    from meta_pde_b200.compat import field, pdes, random as jr
    from meta_pde_b200.compat.flags import FLAGS
    from oracle import meta as ometa
    from oracle.field import nf_apply, unflatten_params
    import helpers
    saved = (FLAGS.bc_weight, FLAGS.xmin)
    FLAGS.bc_weight, FLAGS.xmin = 1.0, -1.0
    try:
        pde = pdes.get_pde("poisson")
        Field = pde.BaseField.partial(sizes=[32, 32], dense_args=(), nonlinearity=field.swish, omega=1.0, omega0=3.0, log_scale=True)
        _, params = Field.init_by_shape(jr.PRNGKey(5), [((1, 2), "float32")])
        model = field.Model(Field, params)
        assert model.net_desc(K).nonlinearity == 1
        prm = pde.sample_params(jr.PRNGKey(1))
        pts = pde.sample_points(jr.PRNGKey(2), 64, prm)
        theta = model.theta.clone().requires_grad_(True)
        b, d = pde.loss_fn(field.Model(model.module, theta), pts, prm)
        loss = torch.stack(list(b.values())).sum() + torch.stack(list(d.values())).sum()
        (g,) = torch.autograd.grad(loss, theta)
        sets = [torch.as_tensor(s, dtype=torch.float32).cpu().reshape(1, -1, 2) for s in pde.abi_sets(pts)]   # oracle side: CPU
        abi_prm = pde.abi_params(prm).reshape(1, -1).cpu()
        lf = ometa.make_loss_fn("poisson", ometa.NetCfg([32, 32], 1.0, True, 1, "swish"), 1.0)
        th64 = model.theta.detach().cpu().double().requires_grad_(True)
        l64, _ = lf(th64, helpers.oracle_points("poisson", sets, 0), helpers.oracle_params("poisson", abi_prm, 0))
        (g64,) = torch.autograd.grad(l64, th64)
        assert rel_l2(loss.detach().cpu(), l64.detach()) < 1e-4 and rel_l2(g.detach().cpu(), g64) < 1e-4
        x = torch.as_tensor(pts[-1], dtype=torch.float32).cpu().reshape(-1, 2)
        u = model(x.to(device))
        want = nf_apply(unflatten_params(th64.detach(), [32, 32], 2, 1, True), 1, x.double(), 1.0, "swish").sum(-1)
        assert rel_l2(u.cpu(), want) < 1e-4
    finally:
        FLAGS.bc_weight, FLAGS.xmin = saved


def test_emu_compat_surface_with_swish(emu_kernels, monkeypatch):
    if torch.cuda.is_available():
        pytest.skip("compat models live on the CUDA device when there is one; the GPU twin below covers it")
    from meta_pde_b200 import ops
    monkeypatch.setattr(ops, "kernels", lambda device=None: emu_kernels)     # TEST ONLY: host logic over the emulator build
    _compat_swish_loss_vs_oracle(emu_kernels, "cpu")


# ------------------------------------------------------------------------------------------- CUDA library
@pytest.mark.gpu
def test_gpu_compat_surface_with_swish(gpu_kernels):
    _compat_swish_loss_vs_oracle(gpu_kernels, "cuda")



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
