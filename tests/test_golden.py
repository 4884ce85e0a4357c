"""Golden fixtures produced by running the reference's own source (tests/golden/make_reference_fixtures.py):
  * not-gpu: the ORACLE must reproduce them in float64 (pins the oracle), and the emulator build of the kernel
    sources must reproduce them at the north-star tolerance (kernel logic, CPU only);
  * gpu: the CUDA kernels through the C ABI must reproduce them at the north-star tolerance."""
import pytest

import golden_cases as gc


def test_fixtures_present():
    assert len(gc.fixtures("loss")) >= 4 and len(gc.fixtures("maml")) >= 4 and len(gc.fixtures("leap")) >= 2


@pytest.mark.parametrize("name", gc.fixtures("loss"))
def test_oracle_loss(name):
    gc.oracle_vs_loss_fixture(name)


@pytest.mark.parametrize("name", gc.fixtures("maml"))
def test_oracle_maml(name):
    gc.oracle_vs_maml_fixture(name)


@pytest.mark.parametrize("name", gc.fixtures("leap"))
def test_oracle_leap(name):
    gc.oracle_vs_leap_fixture(name)


@pytest.mark.parametrize("name", gc.fixtures("loss"))
def test_emu_loss(emu_kernels, name):
    gc.kernels_vs_loss_fixture(emu_kernels, name)


@pytest.mark.parametrize("name", gc.fixtures("maml"))
def test_emu_maml(emu_kernels, name):
    gc.kernels_vs_maml_fixture(emu_kernels, name)


@pytest.mark.parametrize("name", gc.fixtures("leap"))
def test_emu_leap(emu_kernels, name):
    gc.kernels_vs_leap_fixture(emu_kernels, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", gc.fixtures("loss"))
def test_gpu_loss(gpu_kernels, name):
    gc.kernels_vs_loss_fixture(gpu_kernels, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", gc.fixtures("maml"))
def test_gpu_maml(gpu_kernels, name):
    gc.kernels_vs_maml_fixture(gpu_kernels, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", gc.fixtures("leap"))
def test_gpu_leap(gpu_kernels, name):
    gc.kernels_vs_leap_fixture(gpu_kernels, name)
