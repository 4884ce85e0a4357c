"""mpde_field_eval_static: the allocation-free variant of mpde_field_eval (scratch in a file-scope `static __device__`
array of the library, include/metapde_b200.h) against the existing workspace entry point, through the C ABI.

No counterpart in the reference.  The file sorts last on purpose:
a failure here must not hide the parity tests.  It passes alone in a fresh process and after the
other tests (it uses its own tensors and only the shared static array, which every call overwrites before reading)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(K, W, L, out_dim, T, n, seed, shared_theta):
    from meta_pde_b200 import synthetic
    dev = K.device
    theta = synthetic.siren_theta([W] * L, out_dim, 30.0, 30.0, True, seed=seed, bias_scale=0.05).to(dev)
    if not shared_theta:
        g = torch.Generator().manual_seed(seed + 1)
        theta = (theta.unsqueeze(0).repeat(T, 1) + 0.01 * torch.randn(T, theta.numel(), generator=g).to(dev)).contiguous()
    g = torch.Generator().manual_seed(seed + 2)
    coords = (2.0 * torch.rand(T, n, 2, generator=g) - 1.0).to(dev).contiguous()
    return theta, coords


def _static_call(K, net, theta, coords, T, out):
    """Straight through the C ABI; no torch allocation between the pointer hand-over and the launch."""
    P = K.param_count(net)
    st = 0 if theta.dim() == 1 else P
    n = coords.shape[1]
    fn = K._fn("mpde_field_eval_static")
    return fn(C.byref(net), T, C.c_void_p(theta.data_ptr()), st, C.c_void_p(coords.data_ptr()), n * 2, n,
              C.c_void_p(out.data_ptr()), K._stream())


@pytest.mark.parametrize("W,L,out_dim,T,n,shared", [(64, 3, 1, 5, 777, False), (64, 3, 2, 3, 1024, True), (256, 5, 1, 2, 300, False)])
def test_field_eval_static_matches_workspace_entry_point(gpu_kernels, W, L, out_dim, T, n, shared):
    K = gpu_kernels
    net = K.net_desc(L, W, out_dim, 30.0, True)
    assert K.abi.mpde_static_scratch_bytes() == 8 << 20
    # two calls in a row with DIFFERENT inputs: the second must not see anything of the first in the shared array
    for rep, seed in enumerate((3, 11)):
        theta, coords = _case(K, W, L, out_dim, T, n, seed, shared)
        want = K.field_eval(net, theta, coords, T)                      # existing entry point (caller workspace)
        got = torch.full_like(want, float("nan"))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = _static_call(K, net, theta, coords, T, got)
        assert rc == 0, K.abi.mpde_status_string(rc).decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before                  # nothing was allocated through torch
        assert torch.isfinite(got).all()
        # same kernels, same arithmetic, different scratch address: identical; the suite's tolerance is rel-L2 1e-4
        rel = float((got.double() - want.double()).norm() / want.double().norm())
        assert rel <= 1e-4 and torch.equal(got, want), (rep, rel)


def test_field_eval_static_rejects_oversized_problems(gpu_kernels):
    K = gpu_kernels
    net = K.net_desc(5, 256, 1, 30.0, True)
    T = 8                                                               # 8 x 2.2 MB of theta-hat > 8 MiB
    assert K.abi.mpde_field_eval_workspace_bytes(C.byref(net), T) > K.abi.mpde_static_scratch_bytes()
    theta, coords = _case(K, 256, 5, 1, T, 64, 5, False)
    out = torch.zeros(T, 64, 1, device=K.device)
    rc = _static_call(K, net, theta, coords, T, out)
    assert rc == 5 and b"static" in K.abi.mpde_status_string(rc)
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0                                # rejected before any launch
    bad = K.net_desc(3, 48, 1, 30.0, True)
    assert _static_call(K, bad, theta, coords, 1, out) == 2             # MPDE_UNSUPPORTED_NET, as the workspace entry point
