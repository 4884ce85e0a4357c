"""CPU ORACLE for the meta-pde hot path -- TEST INFRASTRUCTURE ONLY.

This package restates, in ``torch.func`` on the CPU, the arithmetic that the
reference (sunnytqin/meta-pde, JAX 0.2.20 + Flax 0.3.3) performs on its hot
path: the SIREN field, the three PDE residual losses, the MAML inner/outer
loop and the LEAP rollout.  It mirrors the reference's autodiff *nesting*
(``vmap(jacfwd(jacrev))``, ``jacfwd(jacfwd)``, ``value_and_grad`` through an
unrolled ``scan``) instead of the hand-derived jets the CUDA kernels use, so
the two are independent derivations of the same numbers.

Who may import this package: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- and there only
as the checker / the timed CPU baseline.  Nothing under ``meta_pde_b200/``
imports it; the product path fails loudly when the CUDA library is missing.

Pinning status: the reference has no tests or golden vectors for this path
 and JAX/Flax are not installable in this image, so the
oracle is pinned by (a) closed-form / finite-difference checks in float64 and
(b) fixtures produced by executing the reference's OWN source files under a
torch-backed stand-in for the ``jax``/``flax`` modules
(``tests/golden/make_reference_fixtures.py``).  See DESIGN.md "Oracle".
"""
