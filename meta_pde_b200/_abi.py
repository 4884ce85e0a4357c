"""ctypes binding of the C ABI in include/metapde_b200.h.

``Abi(path)`` declares every entry point of the shared library; ``Kernels``
wraps them at the tensor level (torch is used for device memory and streams
only).  The product entry point is ``meta_pde_b200.ops.kernels()``, which binds
the nvcc-built ``csrc/libmetapde_b200.so`` and nothing else; tests may bind the
emulator build of the same sources (tests/emu) through this module to check
kernel logic on the CPU-only container.
"""
import ctypes as C
import os

import torch

MPDE_OK = 0
PDE_KINDS = {"poisson": 0, "td_burgers": 1, "hyper_elasticity": 2}
PDE_OUT_DIM = {"poisson": 1, "td_burgers": 1, "hyper_elasticity": 2}
# order of the point sets / loss terms at the ABI (include/metapde_b200.h)
PDE_SETS = {
    "poisson": ["boundary", "domain"],
    "td_burgers": ["left", "right", "initial", "domain"],
    "hyper_elasticity": ["top", "bottom", "domain"],
}
PDE_TERMS = {
    "poisson": ["boundary_loss", "domain_loss"],
    "td_burgers": ["loss_initial", "loss_left", "loss_right", "loss_domain"],
    "hyper_elasticity": ["loss_bottom", "loss_top", "loss_domain"],
}
PDE_NPRM = {"poisson": 13, "td_burgers": 3, "hyper_elasticity": 4}
MODE_LOSS, MODE_GRAD, MODE_HVP = 0, 1, 2


class NetDesc(C.Structure):
    _fields_ = [("num_layers", C.c_int32), ("layer_size", C.c_int32), ("in_dim", C.c_int32),
                ("out_dim", C.c_int32), ("omega", C.c_float), ("log_scale", C.c_int32), ("nonlinearity", C.c_int32)]


NONLINEARITIES = {"sin": 0, "swish": 1}   # mpde_nonlinearity (include/metapde_b200.h)


class PdeDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("bc_weight", C.c_float), ("n_points", C.c_int32 * 4)]


class Points(C.Structure):
    _fields_ = [("set", C.c_void_p * 4), ("task_stride", C.c_int64 * 4)]


class MamlCfg(C.Structure):
    _fields_ = [("inner_steps", C.c_int32), ("inner_lr", C.c_float), ("inner_grad_clip", C.c_float),
                ("softplus_lrs", C.c_int32), ("outer_loss_decay", C.c_float)]


class LeapCfg(C.Structure):
    _fields_ = [("inner_steps", C.c_int32), ("inner_lr", C.c_float), ("inner_grad_clip", C.c_float),
                ("norm", C.c_int32), ("loss_in_distance", C.c_int32), ("stabilize", C.c_int32),
                ("inner_opt", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float)]


# name -> (restype, argtypes); every symbol include/metapde_b200.h declares
_P = C.c_void_p
SIGNATURES = {
    "mpde_param_count": (C.c_int, [C.POINTER(NetDesc)]),
    "mpde_num_point_sets": (C.c_int, [C.c_int]),
    "mpde_num_terms": (C.c_int, [C.c_int]),
    "mpde_num_pde_params": (C.c_int, [C.c_int]),
    "mpde_status_string": (C.c_char_p, [C.c_int]),
    "mpde_workspace_bytes": (C.c_size_t, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.c_int, C.c_int]),
    "mpde_loss": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.c_int, _P, C.c_int64, C.POINTER(Points), _P,
                            _P, _P, _P, _P, C.c_size_t, _P]),
    "mpde_loss_grad": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.c_int, _P, C.c_int64, C.POINTER(Points),
                                 _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "mpde_loss_hvp": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.c_int, _P, C.c_int64, _P, C.c_int64,
                                C.POINTER(Points), _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "mpde_field_eval_workspace_bytes": (C.c_size_t, [C.POINTER(NetDesc), C.c_int]),
    "mpde_field_eval": (C.c_int, [C.POINTER(NetDesc), C.c_int, _P, C.c_int64, _P, C.c_int64, C.c_int, _P, _P,
                                  C.c_size_t, _P]),
    "mpde_random_bits": (C.c_int, [C.c_int, _P, C.c_int64, _P, _P]),
    "mpde_random_uniform": (C.c_int, [C.c_int, _P, C.c_int64, C.c_float, C.c_float, _P, _P]),
    "mpde_random_split": (C.c_int, [C.c_int, _P, C.c_int, _P, _P]),
    "mpde_sample_poisson_candidates": (C.c_int, [C.c_int, _P, C.c_int, _P, C.c_float, C.c_float, C.c_float, C.c_float, _P, _P, _P, _P]),
    "mpde_static_scratch_bytes": (C.c_size_t, []),
    "mpde_field_eval_static": (C.c_int, [C.POINTER(NetDesc), C.c_int, _P, C.c_int64, _P, C.c_int64, C.c_int, _P, _P]),
    "mpde_maml_inner_update": (C.c_int, [C.c_int, C.c_int, _P, C.c_int64, _P, _P, C.c_float, C.c_int, C.c_float,
                                         C.c_float, _P, _P, _P]),
    "mpde_sumsq": (C.c_int, [_P, C.c_int64, _P, _P, C.c_size_t, _P]),
    "mpde_sumsq_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "mpde_adam_step": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, _P,
                                 C.c_float, _P]),
    "mpde_sgd_step": (C.c_int, [_P, _P, C.c_int64, C.c_float, _P, C.c_float, _P]),
    "mpde_task_sum": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_float, _P]),
    "mpde_outer_step": (C.c_int, [_P, C.c_int64, _P, C.c_int64, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_float,
                                  C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                  C.c_float, _P, _P]),
    "mpde_maml_workspace_bytes": (C.c_size_t, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.POINTER(PdeDesc),
                                               C.POINTER(MamlCfg), C.c_int]),
    "mpde_maml_outer": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.POINTER(PdeDesc), C.POINTER(MamlCfg),
                                  C.c_int, _P, _P, C.POINTER(Points), C.POINTER(Points), C.POINTER(Points),
                                  C.POINTER(Points), _P, C.c_float, _P, _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "mpde_maml_adapt": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.POINTER(MamlCfg), C.c_int, _P, _P,
                                  C.POINTER(Points), C.POINTER(Points), _P, _P, _P, _P, C.c_size_t, _P]),
    "mpde_leap_workspace_bytes": (C.c_size_t, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.POINTER(LeapCfg), C.c_int]),
    "mpde_leap_outer": (C.c_int, [C.POINTER(NetDesc), C.POINTER(PdeDesc), C.POINTER(LeapCfg), C.c_int, _P,
                                  C.POINTER(Points), C.POINTER(Points), C.POINTER(Points), _P, C.c_float, _P, _P, _P,
                                  _P, C.c_size_t, _P]),
    "mpde_launch_count": (C.c_int64, []),
    "mpde_backend": (C.c_char_p, []),
    "mpde_prof_enable": (C.c_int, [C.c_int]),
    "mpde_prof_read": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
}


class Abi:
    """The loaded shared library with typed entry points."""

    def __init__(self, path):
        if not os.path.exists(path):
            raise RuntimeError(f"metapde_b200: shared library not found: {path}")
        self.path = path
        self.dll = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(self.dll, name)  # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)

    def backend(self):
        return self.mpde_backend().decode()


class MpdeError(RuntimeError):
    pass


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class Kernels:
    """Tensor-level calls into the ABI.  Every tensor must live on ``device`` and be contiguous float32."""

    def __init__(self, abi, device):
        self.abi = abi
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._ws = None       # grow-only workspace cache (single stream: see workspace())
        self._ws_gen = 0      # bumped whenever the cached workspace is reallocated
        self._sumsq_ws = None
        if self.device.type == "cuda":
            # every ABI entry point is called with this object's device current: the library sets per-device kernel
            # attributes and launches on the caller's stream, which must belong to the current device
            pure = ("mpde_param_count", "mpde_num_point_sets", "mpde_num_terms", "mpde_num_pde_params", "mpde_status_string",
                    "mpde_launch_count", "mpde_backend", "mpde_prof_enable", "mpde_sumsq_workspace_bytes", "mpde_static_scratch_bytes")
            for name in SIGNATURES:
                if name not in pure:       # includes the *_workspace_bytes planners: they query the current device
                    setattr(self, "_" + name, self._guarded(getattr(abi, name)))
        self._fn = lambda name: getattr(self, "_" + name, None) or getattr(self.abi, name)

    def _guarded(self, fn):
        dev = self.device

        def call(*args):
            with torch.cuda.device(dev):
                return fn(*args)
        return call

    # ---------------------------------------------------------------- helpers
    def _check(self, rc):
        if rc != MPDE_OK:
            raise MpdeError(f"metapde_b200: {self.abi.mpde_status_string(rc).decode()} (status {rc})")

    def _stream(self):
        if self.device.type == "cuda":
            return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        return None

    def _t(self, x, name):
        if x is None:
            return None
        if not torch.is_tensor(x) or x.dtype != torch.float32 or x.device.type != self.device.type or not x.is_contiguous():
            raise MpdeError(f"{name}: expected a contiguous float32 tensor on {self.device}")
        if self.device.type == "cuda" and x.device.index != self.device.index:
            raise MpdeError(f"{name}: tensor lives on {x.device}, this Kernels object drives {self.device}")
        return x

    def _eval_bytes(self, net, pd, T, mode):
        nb = self._fn("mpde_workspace_bytes")(C.byref(net), C.byref(pd), T, mode)
        if nb == 0:
            raise MpdeError("metapde_b200: unsupported network / invalid problem description (workspace size 0)")
        return nb

    def _prm(self, pde_params, pde, T):
        prm = self._t(pde_params, "pde_params")
        if prm.shape != (T, PDE_NPRM[pde]):
            raise MpdeError(f"pde_params: expected [T={T}, {PDE_NPRM[pde]}], got {tuple(prm.shape)}")
        return prm

    def workspace(self, nbytes, private=None):
        """The cached grow-only workspace of this object, or ``private`` (a caller-owned uint8 tensor) when given.
        The cached buffer is shared by every call on this object and is therefore SINGLE-STREAM: callers that record
        calls into a CUDA graph or use several streams pass their own ``private`` buffer (MamlTrainer does).  Growing
        frees the old buffer after a device synchronise; ``workspace_generation`` changes when that happens."""
        if private is not None:
            if private.numel() < nbytes:
                raise MpdeError(f"private workspace of {private.numel()} bytes, {nbytes} needed")
            return private
        if self._ws is None or self._ws.numel() < nbytes:
            # release the old buffer first: at C5 size the tensor-core path's spills take 60-115 GB, and old + new
            # would not fit in the 180 GB (the stream is synchronised so that no launch still uses the old one)
            if self._ws is not None and self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
                self._ws = None
                torch.cuda.empty_cache()
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
            self._ws_gen += 1
        return self._ws

    @property
    def workspace_generation(self):
        return self._ws_gen

    def net_desc(self, num_layers, layer_size, out_dim, omega, log_scale=True, in_dim=2, nonlinearity="sin"):
        """nonlinearity: "sin" (SIREN, FLAGS.siren) or "swish" (`--nosiren`, src/maml_pde.py:115; omega is not applied)."""
        if nonlinearity not in NONLINEARITIES:
            raise MpdeError(f"nonlinearity {nonlinearity!r}: expected one of {sorted(NONLINEARITIES)}")
        return NetDesc(int(num_layers), int(layer_size), int(in_dim), int(out_dim), float(omega), int(bool(log_scale)),
                       NONLINEARITIES[nonlinearity])

    def pde_desc(self, pde, bc_weight, n_points):
        n = list(n_points) + [0] * (4 - len(n_points))
        return PdeDesc(PDE_KINDS[pde], float(bc_weight), (C.c_int32 * 4)(*n))

    def param_count(self, net):
        return self.abi.mpde_param_count(C.byref(net))

    def points_struct(self, pde, sets, T):
        """sets: list of tensors [T, n, 2] (or [n, 2]: shared by all tasks) in ABI order -> (Points, n_points)."""
        names = PDE_SETS[pde]
        if len(sets) != len(names):
            raise MpdeError(f"{pde}: expected point sets {names}")
        p = Points()
        n_points = []
        for s, x in enumerate(sets):
            x = self._t(x, f"points[{names[s]}]")
            if x.dim() == 3:
                if x.shape[0] != T or x.shape[2] != 2:
                    raise MpdeError(f"points[{names[s]}]: expected [T={T}, n, 2], got {tuple(x.shape)}")
                p.set[s] = x.data_ptr()
                p.task_stride[s] = x.shape[1] * 2
                n_points.append(x.shape[1])
            elif x.dim() == 2 and x.shape[1] == 2:
                p.set[s] = x.data_ptr()
                p.task_stride[s] = 0
                n_points.append(x.shape[0])
            else:
                raise MpdeError(f"points[{names[s]}]: expected [T, n, 2] or [n, 2], got {tuple(x.shape)}")
        return p, n_points

    def _theta(self, theta, T, P):
        theta = self._t(theta, "theta")
        if theta.dim() == 1 and theta.shape[0] == P:
            return theta, 0
        if theta.dim() == 2 and theta.shape == (T, P):
            return theta, P
        raise MpdeError(f"theta: expected [P={P}] or [T={T}, P={P}], got {tuple(theta.shape)}")

    # ---------------------------------------------------------------- K1-K3
    def loss(self, net, pde, bc_weight, theta, sets, pde_params, T, want_residual=False):
        P = self.param_count(net)
        theta, st = self._theta(theta, T, P)
        pts, n_points = self.points_struct(pde, sets, T)
        pd = self.pde_desc(pde, bc_weight, n_points)
        prm = self._prm(pde_params, pde, T)
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        res = torch.empty(T, n_points[-1], dtype=torch.float32, device=self.device) if want_residual else None
        ws = self.workspace(self._eval_bytes(net, pd, T, MODE_LOSS))
        self._check(self._fn("mpde_loss")(C.byref(net), C.byref(pd), T, _ptr(theta), st, C.byref(pts), _ptr(prm),
                                       _ptr(loss), _ptr(terms), _ptr(res), _ptr(ws), ws.numel(), self._stream()))
        return (loss, terms, res) if want_residual else (loss, terms)

    def loss_grad(self, net, pde, bc_weight, theta, sets, pde_params, T):
        P = self.param_count(net)
        theta, st = self._theta(theta, T, P)
        pts, n_points = self.points_struct(pde, sets, T)
        pd = self.pde_desc(pde, bc_weight, n_points)
        prm = self._prm(pde_params, pde, T)
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        grad = torch.empty(T, P, dtype=torch.float32, device=self.device)
        ws = self.workspace(self._eval_bytes(net, pd, T, MODE_GRAD))
        self._check(self._fn("mpde_loss_grad")(C.byref(net), C.byref(pd), T, _ptr(theta), st, C.byref(pts), _ptr(prm),
                                            _ptr(loss), _ptr(terms), _ptr(grad), _ptr(ws), ws.numel(), self._stream()))
        return loss, terms, grad

    def loss_hvp(self, net, pde, bc_weight, theta, v, sets, pde_params, T, want_grad=True):
        P = self.param_count(net)
        theta, st = self._theta(theta, T, P)
        v = self._t(v, "v")
        if v.dim() == 1 and v.shape[0] == P:
            vst = 0
        elif v.shape == (T, P):
            vst = P
        else:
            raise MpdeError(f"v: expected [P] or [T, P], got {tuple(v.shape)}")
        pts, n_points = self.points_struct(pde, sets, T)
        pd = self.pde_desc(pde, bc_weight, n_points)
        prm = self._prm(pde_params, pde, T)
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        grad = torch.empty(T, P, dtype=torch.float32, device=self.device) if want_grad else None
        hv = torch.empty(T, P, dtype=torch.float32, device=self.device)
        ws = self.workspace(self._eval_bytes(net, pd, T, MODE_HVP))
        self._check(self._fn("mpde_loss_hvp")(C.byref(net), C.byref(pd), T, _ptr(theta), st, _ptr(v), vst, C.byref(pts),
                                           _ptr(prm), _ptr(loss), _ptr(terms), _ptr(grad), _ptr(hv), _ptr(ws),
                                           ws.numel(), self._stream()))
        return loss, terms, grad, hv

    def field_eval(self, net, theta, coords, T):
        P = self.param_count(net)
        theta, st = self._theta(theta, T, P)
        coords = self._t(coords, "coords")
        if coords.dim() == 3:
            n, cst = coords.shape[1], coords.shape[1] * 2
        else:
            n, cst = coords.shape[0], 0
        out = torch.empty(T, n, net.out_dim, dtype=torch.float32, device=self.device)
        nb = self._fn("mpde_field_eval_workspace_bytes")(C.byref(net), T)
        ws = self.workspace(max(nb, 1))
        self._check(self._fn("mpde_field_eval")(C.byref(net), T, _ptr(theta), st, _ptr(coords), cst, n, _ptr(out),
                                             _ptr(ws), ws.numel(), self._stream()))
        return out

    # ---------------------------------------------------------------- K4
    def maml_inner_update(self, theta, grad, alpha, softplus_lrs, inner_lr, clip):
        T, P = grad.shape
        theta, st = self._theta(theta, T, P)
        out = torch.empty(T, P, dtype=torch.float32, device=self.device)
        gn = torch.empty(T, dtype=torch.float32, device=self.device)
        a_t = alpha if torch.is_tensor(alpha) else None
        a_s = 1.0 if torch.is_tensor(alpha) else float(alpha)
        self._check(self._fn("mpde_maml_inner_update")(T, P, _ptr(theta), st, _ptr(self._t(grad, "grad")),
                                                    _ptr(self._t(a_t, "alpha")), a_s, int(softplus_lrs),
                                                    float(inner_lr), float(clip), _ptr(out), _ptr(gn), self._stream()))
        return out, gn

    # ---------------------------------------------------------------- samplers (jax.random on the GPU)
    def _keys(self, keys):
        """[S, 2] uint32 key pairs (numpy / tensor) -> int32 device tensor holding the same bits"""
        import numpy as _np
        if not torch.is_tensor(keys):
            keys = torch.from_numpy(_np.ascontiguousarray(_np.asarray(keys, dtype=_np.uint32).reshape(-1, 2)).view(_np.int32))
        if keys.dtype != torch.int32 or keys.dim() != 2 or keys.shape[1] != 2:
            raise MpdeError("keys: expected [S, 2] uint32 key pairs")
        return keys.to(self.device).contiguous()

    def random_bits(self, keys, n):
        k = self._keys(keys)
        out = torch.empty(k.shape[0], int(n), dtype=torch.int32, device=self.device)
        self._check(self._fn("mpde_random_bits")(k.shape[0], _ptr(k), int(n), _ptr(out), self._stream()))
        return out

    def random_uniform(self, keys, n, minval=0.0, maxval=1.0):
        k = self._keys(keys)
        out = torch.empty(k.shape[0], int(n), dtype=torch.float32, device=self.device)
        self._check(self._fn("mpde_random_uniform")(k.shape[0], _ptr(k), int(n), float(minval), float(maxval), _ptr(out), self._stream()))
        return out

    def random_split(self, keys, num):
        k = self._keys(keys)
        out = torch.empty(k.shape[0], int(num), 2, dtype=torch.int32, device=self.device)
        self._check(self._fn("mpde_random_split")(k.shape[0], _ptr(k), int(num), _ptr(out), self._stream()))
        return out

    def sample_poisson_candidates(self, keys, n, geo, xmin, xmax, ymin, ymax):
        k = self._keys(keys)
        S, n = k.shape[0], int(n)
        geo = self._t(torch.as_tensor(geo, dtype=torch.float32).to(self.device).reshape(S, 2).contiguous(), "geo")
        boundary = torch.empty(S, n, 2, dtype=torch.float32, device=self.device)
        cand_xy = torch.empty(S, 3 * n, 2, dtype=torch.float32, device=self.device)
        cand_g = torch.empty(S, 3 * n, dtype=torch.float32, device=self.device)
        self._check(self._fn("mpde_sample_poisson_candidates")(S, _ptr(k), n, _ptr(geo), float(xmin), float(xmax), float(ymin),
                                                              float(ymax), _ptr(boundary), _ptr(cand_xy), _ptr(cand_g), self._stream()))
        return boundary, cand_xy, cand_g

    def prof_enable(self, on=True):
        self._check(self.abi.mpde_prof_enable(int(bool(on))))

    def prof_read(self, mode):
        """-> (total_ms, launches) of the jet-kernel launches of ``mode`` since the last read."""
        ms, n = C.c_double(0.0), C.c_int64(0)
        self._check(self._fn("mpde_prof_read")(int(mode), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        return int(self.abi.mpde_launch_count())

    def sumsq(self, x):
        out = torch.empty(1, dtype=torch.float32, device=self.device)
        nb = self.abi.mpde_sumsq_workspace_bytes(x.numel())
        if self._sumsq_ws is None or self._sumsq_ws.numel() < nb:
            self._sumsq_ws = torch.empty(int(nb), dtype=torch.uint8, device=self.device)
        ws = self._sumsq_ws
        self._check(self._fn("mpde_sumsq")(_ptr(self._t(x, "x")), x.numel(), _ptr(out), _ptr(ws), ws.numel(), self._stream()))
        return out

    def adam_step(self, x, g, m, v, t, lr, b1, b2, eps=1e-8, sumsq=None, clip=0.0):
        self._check(self._fn("mpde_adam_step")(_ptr(self._t(x, "x")), _ptr(self._t(g, "g")), _ptr(self._t(m, "m")),
                                            _ptr(self._t(v, "v")), x.numel(), int(t), float(lr), float(b1), float(b2),
                                            float(eps), _ptr(sumsq), float(clip), self._stream()))

    def sgd_step(self, x, g, lr, sumsq=None, clip=0.0):
        self._check(self._fn("mpde_sgd_step")(_ptr(self._t(x, "x")), _ptr(self._t(g, "g")), x.numel(), float(lr),
                                           _ptr(sumsq), float(clip), self._stream()))

    def task_sum(self, x, scale=1.0, out=None):
        """[T, P] -> [P]: scale * sum over tasks, fixed order (maml.py:231 with scale = 1/T)."""
        T, P = x.shape
        out = torch.empty(P, dtype=torch.float32, device=self.device) if out is None else out
        self._check(self._fn("mpde_task_sum")(_ptr(self._t(out, "out")), _ptr(self._t(x, "x")), int(T), int(P), float(scale),
                                              self._stream()))
        return out

    def outer_step(self, theta, g_packed, m_theta, v_theta, t_theta, lr_theta, b1, b2, eps=1e-8, clip=1e14,
                   optimizer="adam", lrs=None, m_lrs=None, v_lrs=None, t_lrs=1, lr_lrs=0.0, b1_lrs=0.9, b2_lrs=0.99,
                   eps_lrs=1e-8, out=None):
        """One launch: ||g_theta||, clip of [g_theta | g_lrs] by it, optimizer step on theta, jax-adam step on the lr
        tree (src/maml_pde.py:210-225).  Returns the device scalar sum(g_theta^2)."""
        n_theta = theta.numel()
        n_lrs = 0 if lrs is None else lrs.numel()
        if g_packed.numel() != n_theta + n_lrs:
            raise MpdeError(f"g_packed: expected {n_theta + n_lrs} floats, got {g_packed.numel()}")
        out = torch.empty(1, dtype=torch.float32, device=self.device) if out is None else out
        adam = optimizer != "sgd"
        self._check(self._fn("mpde_outer_step")(
            _ptr(self._t(theta, "theta")), n_theta, _ptr(self._t(lrs, "lrs")), n_lrs, _ptr(self._t(g_packed, "g_packed")),
            _ptr(self._t(m_theta, "m_theta") if adam else None), _ptr(self._t(v_theta, "v_theta") if adam else None),
            _ptr(self._t(m_lrs, "m_lrs")), _ptr(self._t(v_lrs, "v_lrs")), 1 if adam else 0, int(t_theta), float(lr_theta),
            float(b1), float(b2), float(eps), int(t_lrs), float(lr_lrs), float(b1_lrs), float(b2_lrs), float(eps_lrs),
            float(clip), _ptr(out), self._stream()))
        return out

    # ---------------------------------------------------------------- rollouts
    def _points_array(self, pde, list_of_sets, T):
        arr = (Points * len(list_of_sets))()
        n_points = None
        for k, sets in enumerate(list_of_sets):
            p, n = self.points_struct(pde, sets, T)
            arr[k] = p
            if n_points is None:
                n_points = n
            elif n != n_points:
                raise MpdeError("all inner steps must use the same number of points per set")
        return arr, n_points

    def maml_workspace_bytes(self, net, pde, bc_weight, cfg, n_inner, n_outer, T):
        mc = MamlCfg(int(cfg["inner_steps"]), float(cfg["inner_lr"]), float(cfg["inner_grad_clip"]),
                     int(bool(cfg["softplus_lrs"])), float(cfg["outer_loss_decay"]))
        pdi, pdo = self.pde_desc(pde, bc_weight, n_inner), self.pde_desc(pde, bc_weight, n_outer)
        return self._fn("mpde_maml_workspace_bytes")(C.byref(net), C.byref(pdi), C.byref(pdo), C.byref(mc), T)

    def maml_outer(self, net, pde, bc_weight, cfg, theta0, inner_lrs, inner_sets, outer_sets, final_sets, aux_sets,
                   pde_params, T, grad_scale=None, want_theta_final=False, out_g_theta=None, out_g_lrs=None,
                   workspace=None):
        """cfg: dict(inner_steps, inner_lr, inner_grad_clip, softplus_lrs, outer_loss_decay).
        inner_sets / outer_sets: K lists of point tensors (ABI order); final_sets / aux_sets: one list each."""
        P = self.param_count(net)
        K = int(cfg["inner_steps"])
        theta0 = self._t(theta0, "theta0")
        mc = MamlCfg(K, float(cfg["inner_lr"]), float(cfg["inner_grad_clip"]), int(bool(cfg["softplus_lrs"])),
                     float(cfg["outer_loss_decay"]))
        ipts, n_in = self._points_array(pde, inner_sets, T)
        opts, n_out = self._points_array(pde, outer_sets, T)
        fpts, n_f = self.points_struct(pde, final_sets, T)
        apts, n_a = self.points_struct(pde, aux_sets, T)
        if n_f != n_in or n_a != n_out:
            raise MpdeError("final points must match the inner sizes and aux points the outer sizes")
        pdi = self.pde_desc(pde, bc_weight, n_in)
        pdo = self.pde_desc(pde, bc_weight, n_out)
        prm = self._prm(pde_params, pde, T)
        lrs = self._t(inner_lrs, "inner_lrs")
        nt = len(PDE_TERMS[pde])
        dev = self.device
        g_theta = self._t(out_g_theta, "out_g_theta") if out_g_theta is not None else torch.empty(P, dtype=torch.float32, device=dev)
        if g_theta.numel() != P:
            raise MpdeError(f"out_g_theta: expected {P} floats")
        g_lrs = None
        if lrs is not None:
            g_lrs = self._t(out_g_lrs, "out_g_lrs") if out_g_lrs is not None else torch.empty(K, P, dtype=torch.float32, device=dev)
            if g_lrs.numel() != K * P:
                raise MpdeError(f"out_g_lrs: expected [{K}, {P}]")
        losses = torch.empty(T, K + 1, dtype=torch.float32, device=dev)
        meta = torch.empty(T, dtype=torch.float32, device=dev)
        aux = torch.empty(T, nt, dtype=torch.float32, device=dev)
        thf = torch.empty(T, P, dtype=torch.float32, device=dev) if want_theta_final else None
        nb = self._fn("mpde_maml_workspace_bytes")(C.byref(net), C.byref(pdi), C.byref(pdo), C.byref(mc), T)
        if nb == 0:
            raise MpdeError("metapde_b200: unsupported network / invalid MAML configuration (workspace size 0)")
        ws = self.workspace(nb, private=workspace)
        gs = (1.0 / T) if grad_scale is None else float(grad_scale)
        self._check(self._fn("mpde_maml_outer")(C.byref(net), C.byref(pdi), C.byref(pdo), C.byref(mc), T, _ptr(theta0),
                                             _ptr(lrs), ipts, opts, C.byref(fpts), C.byref(apts), _ptr(prm), gs,
                                             _ptr(g_theta), _ptr(g_lrs), _ptr(losses), _ptr(meta), _ptr(aux), _ptr(thf),
                                             _ptr(ws), ws.numel(), self._stream()))
        return {"g_theta": g_theta, "g_lrs": g_lrs, "losses": losses, "meta_losses": meta, "aux_terms": aux,
                "theta_final": thf}

    def maml_adapt(self, net, pde, bc_weight, cfg, theta0, inner_lrs, inner_sets, final_sets, pde_params, T):
        P = self.param_count(net)
        K = int(cfg["inner_steps"])
        mc = MamlCfg(K, float(cfg["inner_lr"]), float(cfg["inner_grad_clip"]), int(bool(cfg["softplus_lrs"])),
                     float(cfg.get("outer_loss_decay", 0.0)))
        ipts, n_in = self._points_array(pde, inner_sets, T)
        fpts, n_f = self.points_struct(pde, final_sets, T)
        pdi = self.pde_desc(pde, bc_weight, n_in)
        prm = self._prm(pde_params, pde, T)
        lrs = self._t(inner_lrs, "inner_lrs")
        thf = torch.empty(T, P, dtype=torch.float32, device=self.device)
        losses = torch.empty(T, K + 1, dtype=torch.float32, device=self.device)
        nb = self._fn("mpde_maml_workspace_bytes")(C.byref(net), C.byref(pdi), None, C.byref(mc), T)
        ws = self.workspace(max(nb, 1))
        self._check(self._fn("mpde_maml_adapt")(C.byref(net), C.byref(pdi), C.byref(mc), T, _ptr(self._t(theta0, "theta0")),
                                             _ptr(lrs), ipts, C.byref(fpts), _ptr(prm), _ptr(thf), _ptr(losses),
                                             _ptr(ws), ws.numel(), self._stream()))
        return thf, losses

    def leap_outer(self, net, pde, bc_weight, cfg, theta0, loss0_sets, step_sets, new_sets, pde_params, T,
                   grad_scale=None, want_theta_final=False, out_meta_grad=None):
        """cfg: dict(inner_steps, inner_lr, inner_grad_clip, norm, loss_in_distance, stabilize, inner_opt, beta1, beta2, eps)."""
        P = self.param_count(net)
        K = int(cfg["inner_steps"])
        lc = LeapCfg(K, float(cfg["inner_lr"]), float(cfg["inner_grad_clip"]), int(bool(cfg.get("norm", True))),
                     int(bool(cfg.get("loss_in_distance", True))), int(bool(cfg.get("stabilize", True))),
                     1 if cfg.get("inner_opt", "adam") == "adam" else 0, float(cfg.get("beta1", 0.9)),
                     float(cfg.get("beta2", 0.99)), float(cfg.get("eps", 1e-8)))
        l0, n0 = self.points_struct(pde, loss0_sets, T)
        sp, n1 = self._points_array(pde, step_sets, T)
        npn, n2 = self._points_array(pde, new_sets, T)
        if not (n0 == n1 == n2):
            raise MpdeError("LEAP point sets must all have the same sizes")
        pd = self.pde_desc(pde, bc_weight, n0)
        prm = self._prm(pde_params, pde, T)
        g = self._t(out_meta_grad, "out_meta_grad") if out_meta_grad is not None else torch.empty(P, dtype=torch.float32, device=self.device)
        if g.numel() != P:
            raise MpdeError(f"out_meta_grad: expected {P} floats")
        losses = torch.empty(T, K + 1, dtype=torch.float32, device=self.device)
        thf = torch.empty(T, P, dtype=torch.float32, device=self.device) if want_theta_final else None
        nb = self._fn("mpde_leap_workspace_bytes")(C.byref(net), C.byref(pd), C.byref(lc), T)
        ws = self.workspace(max(nb, 1))
        gs = (1.0 / T) if grad_scale is None else float(grad_scale)
        self._check(self._fn("mpde_leap_outer")(C.byref(net), C.byref(pd), C.byref(lc), T, _ptr(self._t(theta0, "theta0")),
                                             C.byref(l0), sp, npn, _ptr(prm), gs, _ptr(g), _ptr(losses), _ptr(thf),
                                             _ptr(ws), ws.numel(), self._stream()))
        return {"meta_grad": g, "losses": losses, "theta_final": thf}
