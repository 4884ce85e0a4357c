"""Host-side mirror of the reference's plugin surface, dispatching to the sm_100a kernels.

Same names, argument meaning and error behaviour as the reference (paths relative to /root/reference):
    compat.get_pde.get_pde(name)             src/get_pde.py:10-24  -> per-PDE module objects with sample_params,
                                             sample_points, sample_points_in_domain, loss_fn, BaseField, dim
    compat.field                             src/nets/field.py     (NeuralPotential, make_nf_ndim, Model, SIREN init)
    compat.maml / compat.leap                src/nets/maml.py, src/nets/leap.py (MamlDef / LeapDef, rollouts, meta-grads)
    compat.flags                             src/util/common_flags.py + the drivers' own flags
    compat.maml_pde / leap_pde / nn_pde_maml src/maml_pde.py, src/leap_pde.py, src/nn_pde_maml.py (absl apps)
    compat.random                            jax.random (threefry2x32) for the samplers
Arrays are torch CUDA tensors instead of jax DeviceArrays; models are pytrees over ONE flat theta tensor.
There is no CPU fallback: anything that needs arithmetic on the path calls libmetapde_b200.so.
"""
