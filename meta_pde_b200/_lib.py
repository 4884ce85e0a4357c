"""Loader of the product library csrc/libmetapde_b200.so (nvcc, sm_100a).

Fails loudly: there is no CPU path.  The emulator build under tests/emu is
test infrastructure and is never looked for here.
"""
import os
import threading

from ._abi import Abi, Kernels

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libmetapde_b200.so")
_lock = threading.Lock()
_abi = None
_kernels = {}


def lib_path():
    return _LIB_PATH


def abi():
    """The loaded C ABI (symbols only; usable without a GPU for introspection)."""
    global _abi
    with _lock:
        if _abi is None:
            if not os.path.exists(_LIB_PATH):
                raise RuntimeError(
                    f"metapde_b200: {_LIB_PATH} is missing. Build it with "
                    "`python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
                    "There is no CPU fallback.")
            a = Abi(_LIB_PATH)
            if a.backend() != "cuda":
                raise RuntimeError(f"metapde_b200: {_LIB_PATH} is not the CUDA build")
            _abi = a
        return _abi


def kernels(device=None):
    """Tensor-level kernel interface bound to a CUDA device.  Raises without CUDA."""
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("metapde_b200: CUDA is not available; this package has no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError(f"metapde_b200: device {dev} is not a CUDA device; there is no CPU fallback")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    a = abi()
    with _lock:
        if dev not in _kernels:
            _kernels[dev] = Kernels(a, dev)
        return _kernels[dev]
