"""Public tensor-level ops of the hot path (thin names over ``_lib.kernels()``)."""
from ._lib import abi, kernels, lib_path  # noqa: F401
from ._abi import MpdeError, PDE_KINDS, PDE_NPRM, PDE_OUT_DIM, PDE_SETS, PDE_TERMS  # noqa: F401
