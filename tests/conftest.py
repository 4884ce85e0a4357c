import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def emu_kernels():
    """Kernels bound to the emulator build of the kernel sources (test infrastructure)."""
    import subprocess

    from meta_pde_b200._abi import Abi, Kernels

    d = os.path.join(ROOT, "tests", "emu")
    subprocess.run(["make", "-C", d, "-j", "8"], check=True, capture_output=True)
    return Kernels(Abi(os.path.join(d, "libmetapde_emu.so")), "cpu")


@pytest.fixture(scope="session")
def gpu_kernels():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from meta_pde_b200 import ops

    return ops.kernels(torch.device("cuda", 0))
