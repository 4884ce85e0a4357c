"""meta_pde_b200 -- B200-native (sm_100a) implementation of meta-pde's data-parallel hot path.

Hand-written CUDA kernels behind the C ABI of include/metapde_b200.h
(csrc/libmetapde_b200.so), plus the host-side mirror of the reference's plugin
surface (get_pde, per-PDE modules, nets.field / nets.maml / nets.leap, drivers).
There is no CPU fallback: every compute entry point raises if the CUDA library
is missing or the tensors are not on a CUDA device.
"""
__version__ = "0.1.0"
