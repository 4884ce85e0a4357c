"""CPU-only checks of the kernel LOGIC: the kernel sources of meta_pde_b200/csrc are compiled for the
host-side SIMT emulator (tests/emu, test infrastructure) and driven through the same C ABI and the same
checkers as the GPU parity tests.  Small sizes so the whole CPU suite stays within minutes."""
import pytest

import parity_cases as pc


@pytest.mark.parametrize("pde,W,L,n", [
    ("poisson", 32, 2, 70),          # ragged tiles (70 = 2 full tiles + 6 points)
    ("td_burgers", 32, 2, 70),
    ("hyper_elasticity", 32, 2, 50),
    ("poisson", 64, 3, 40),
    ("poisson", 32, 1, 33),          # single hidden layer: no hidden-hidden contraction
])
def test_loss_grad_hvp(emu_kernels, pde, W, L, n):
    pc.check_loss_grad_hvp(emu_kernels, pde, W, L, n, T=2)


def test_shared_theta_and_bc_weight(emu_kernels):
    pc.check_loss_grad_hvp(emu_kernels, "poisson", 32, 2, 40, T=3, bc_weight=100.0, per_task_theta=False)


def test_wide_256_tile(emu_kernels):
    # the W=256 geometry (16-point tiles, 4 column passes in the weight-gradient contraction)
    pc.check_loss_grad_hvp(emu_kernels, "td_burgers", 256, 2, 20, T=1, set_sizes=[5, 5, 22, 20])


def test_field_eval(emu_kernels):
    pc.check_field_eval(emu_kernels, 32, 2, 1, 300, T=2)
    pc.check_field_eval(emu_kernels, 32, 2, 2, 130, T=1)


@pytest.mark.parametrize("pde", ["poisson", "td_burgers", "hyper_elasticity"])
def test_maml_outer(emu_kernels, pde):
    lr = 1e-2 if pde != "hyper_elasticity" else 1e-4
    pc.check_maml(emu_kernels, pde, 32, 2, 40, T=2, K_steps=2, inner_lr=lr)


def test_maml_outer_clip_active(emu_kernels):
    # clip far below the gradient norm: the VJP of the normalisation branch (maml.py:87-94)
    pc.check_maml(emu_kernels, "poisson", 32, 2, 40, T=2, K_steps=3, inner_lr=1e-2, clip=1e-2, decay=0.5)


def test_maml_outer_scalar_lr(emu_kernels):
    pc.check_maml(emu_kernels, "poisson", 32, 2, 40, T=2, K_steps=2, inner_lr=1e-2, lr_tree=False)


def test_maml_adapt(emu_kernels):
    pc.check_maml_adapt(emu_kernels, "poisson", 32, 2, 40, T=2, K_steps=3, inner_lr=1e-2)


@pytest.mark.parametrize("opt", ["adam", "sgd"])
def test_leap_outer(emu_kernels, opt):
    pc.check_leap(emu_kernels, "poisson", 32, 2, 40, T=2, K_steps=3, inner_lr=1e-3 if opt == "adam" else 1e-2, inner_opt=opt)


def test_leap_outer_clip(emu_kernels):
    pc.check_leap(emu_kernels, "td_burgers", 32, 2, 70, T=2, K_steps=2, inner_lr=1e-3, clip=0.5)


# ---- outer train steps vs oracle/meta.py (rows a13, a14-outer, a15, f1): fused clip + Adam(theta) + Adam(alpha) kernel ----
@pytest.mark.parametrize("grad_clip", [1e14, 1e-3])
def test_maml_train_step_matches_oracle(emu_kernels, grad_clip):
    errs, clipped = pc.check_maml_train_steps(emu_kernels, "poisson", 32, 2, 30, T=2, K_steps=2, steps=3, grad_clip=grad_clip)
    assert all(clipped) == (grad_clip < 1.0) and any(clipped) == (grad_clip < 1.0), clipped


def test_maml_train_step_sgd_and_rmsprop(emu_kernels):
    pc.check_maml_train_steps(emu_kernels, "poisson", 32, 2, 30, T=2, K_steps=2, steps=2, optimizer="sgd")
    pc.check_maml_train_steps(emu_kernels, "poisson", 32, 2, 30, T=2, K_steps=2, steps=2, optimizer="rmsprop")


@pytest.mark.parametrize("grad_clip", [1e14, 1e-3])
def test_leap_train_step_matches_oracle(emu_kernels, grad_clip):
    errs, clipped = pc.check_leap_train_steps(emu_kernels, "poisson", 32, 2, 30, T=2, K_steps=2, steps=3, grad_clip=grad_clip)
    assert all(clipped) == (grad_clip < 1.0), clipped


@pytest.mark.parametrize("grad_clip", [1e14, 1e-3])
def test_finetune_train_step_matches_oracle(emu_kernels, grad_clip):
    errs, clipped = pc.check_finetune_train_steps(emu_kernels, "td_burgers", 32, 2, 70, T=2, steps=3, grad_clip=grad_clip)
    assert all(clipped) == (grad_clip < 1.0), clipped


def test_gpu_sampler_kernels_on_the_emulator(emu_kernels):
    """jax.random (threefry split / bits / uniform) and the batched Poisson / Burgers point samplers of
    csrc/sampler_kernels.cuh, run through the emulator build: bit-exact integers and uniforms, same points as the host
    samplers that tests/test_sampler_parity.py pins to the reference's sampler source."""
    import sampler_cases as sc
    sc.check_threefry(emu_kernels)
    sc.check_poisson(emu_kernels)
    sc.check_burgers(emu_kernels)
