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
                ("out_dim", C.c_int32), ("omega", C.c_float), ("log_scale", C.c_int32)]


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
    "mpde_maml_inner_update": (C.c_int, [C.c_int, C.c_int, _P, C.c_int64, _P, _P, C.c_float, C.c_int, C.c_float,
                                         C.c_float, _P, _P, _P]),
    "mpde_sumsq": (C.c_int, [_P, C.c_int64, _P, _P, C.c_size_t, _P]),
    "mpde_sumsq_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "mpde_adam_step": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, _P,
                                 C.c_float, _P]),
    "mpde_sgd_step": (C.c_int, [_P, _P, C.c_int64, C.c_float, _P, C.c_float, _P]),
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
        self._ws = None  # grow-only workspace cache

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
        return x

    def workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            # release the old buffer first: at C5 size the tensor-core path's spills take 60-115 GB, and old + new
            # would not fit in the 180 GB (the stream is synchronised so that no launch still uses the old one)
            if self._ws is not None and self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
                self._ws = None
                torch.cuda.empty_cache()
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self._ws

    def net_desc(self, num_layers, layer_size, out_dim, omega, log_scale=True, in_dim=2):
        return NetDesc(int(num_layers), int(layer_size), int(in_dim), int(out_dim), float(omega), int(bool(log_scale)))

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
        prm = self._t(pde_params, "pde_params")
        if prm.shape != (T, PDE_NPRM[pde]):
            raise MpdeError(f"pde_params: expected [T={T}, {PDE_NPRM[pde]}], got {tuple(prm.shape)}")
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        res = torch.empty(T, n_points[-1], dtype=torch.float32, device=self.device) if want_residual else None
        nb = self.abi.mpde_workspace_bytes(C.byref(net), C.byref(pd), T, MODE_LOSS)
        if nb == 0:
            self._check(self.abi.mpde_loss(C.byref(net), C.byref(pd), T, _ptr(theta), st, C.byref(pts), _ptr(prm),
                                           _ptr(loss), _ptr(terms), _ptr(res), None, 0, self._stream()))
        ws = self.workspace(nb)
        self._check(self.abi.mpde_loss(C.byref(net), C.byref(pd), T, _ptr(theta), st, C.byref(pts), _ptr(prm),
                                       _ptr(loss), _ptr(terms), _ptr(res), _ptr(ws), ws.numel(), self._stream()))
        return (loss, terms, res) if want_residual else (loss, terms)

    def loss_grad(self, net, pde, bc_weight, theta, sets, pde_params, T):
        P = self.param_count(net)
        theta, st = self._theta(theta, T, P)
        pts, n_points = self.points_struct(pde, sets, T)
        pd = self.pde_desc(pde, bc_weight, n_points)
        prm = self._t(pde_params, "pde_params")
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        grad = torch.empty(T, P, dtype=torch.float32, device=self.device)
        nb = self.abi.mpde_workspace_bytes(C.byref(net), C.byref(pd), T, MODE_GRAD)
        ws = self.workspace(max(nb, 1))
        self._check(self.abi.mpde_loss_grad(C.byref(net), C.byref(pd), T, _ptr(theta), st, C.byref(pts), _ptr(prm),
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
        prm = self._t(pde_params, "pde_params")
        nt = len(PDE_TERMS[pde])
        loss = torch.empty(T, dtype=torch.float32, device=self.device)
        terms = torch.empty(T, nt, dtype=torch.float32, device=self.device)
        grad = torch.empty(T, P, dtype=torch.float32, device=self.device) if want_grad else None
        hv = torch.empty(T, P, dtype=torch.float32, device=self.device)
        nb = self.abi.mpde_workspace_bytes(C.byref(net), C.byref(pd), T, MODE_HVP)
        ws = self.workspace(max(nb, 1))
        self._check(self.abi.mpde_loss_hvp(C.byref(net), C.byref(pd), T, _ptr(theta), st, _ptr(v), vst, C.byref(pts),
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
        nb = self.abi.mpde_field_eval_workspace_bytes(C.byref(net), T)
        ws = self.workspace(max(nb, 1))
        self._check(self.abi.mpde_field_eval(C.byref(net), T, _ptr(theta), st, _ptr(coords), cst, n, _ptr(out),
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
        self._check(self.abi.mpde_maml_inner_update(T, P, _ptr(theta), st, _ptr(self._t(grad, "grad")),
                                                    _ptr(self._t(a_t, "alpha")), a_s, int(softplus_lrs),
                                                    float(inner_lr), float(clip), _ptr(out), _ptr(gn), self._stream()))
        return out, gn

    def prof_enable(self, on=True):
        self._check(self.abi.mpde_prof_enable(int(bool(on))))

    def prof_read(self, mode):
        """-> (total_ms, launches) of the jet-kernel launches of ``mode`` since the last read."""
        ms, n = C.c_double(0.0), C.c_int64(0)
        self._check(self.abi.mpde_prof_read(int(mode), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        return int(self.abi.mpde_launch_count())

    def sumsq(self, x):
        out = torch.empty(1, dtype=torch.float32, device=self.device)
        nb = self.abi.mpde_sumsq_workspace_bytes(x.numel())
        ws = torch.empty(int(nb), dtype=torch.uint8, device=self.device)
        self._check(self.abi.mpde_sumsq(_ptr(self._t(x, "x")), x.numel(), _ptr(out), _ptr(ws), ws.numel(), self._stream()))
        return out

    def adam_step(self, x, g, m, v, t, lr, b1, b2, eps=1e-8, sumsq=None, clip=0.0):
        self._check(self.abi.mpde_adam_step(_ptr(self._t(x, "x")), _ptr(self._t(g, "g")), _ptr(self._t(m, "m")),
                                            _ptr(self._t(v, "v")), x.numel(), int(t), float(lr), float(b1), float(b2),
                                            float(eps), _ptr(sumsq), float(clip), self._stream()))

    def sgd_step(self, x, g, lr, sumsq=None, clip=0.0):
        self._check(self.abi.mpde_sgd_step(_ptr(self._t(x, "x")), _ptr(self._t(g, "g")), x.numel(), float(lr),
                                           _ptr(sumsq), float(clip), self._stream()))

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

    def maml_outer(self, net, pde, bc_weight, cfg, theta0, inner_lrs, inner_sets, outer_sets, final_sets, aux_sets,
                   pde_params, T, grad_scale=None, want_theta_final=False, out_g_theta=None, out_g_lrs=None):
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
        prm = self._t(pde_params, "pde_params")
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
        nb = self.abi.mpde_maml_workspace_bytes(C.byref(net), C.byref(pdi), C.byref(pdo), C.byref(mc), T)
        ws = self.workspace(max(nb, 1))
        gs = (1.0 / T) if grad_scale is None else float(grad_scale)
        self._check(self.abi.mpde_maml_outer(C.byref(net), C.byref(pdi), C.byref(pdo), C.byref(mc), T, _ptr(theta0),
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
        prm = self._t(pde_params, "pde_params")
        lrs = self._t(inner_lrs, "inner_lrs")
        thf = torch.empty(T, P, dtype=torch.float32, device=self.device)
        losses = torch.empty(T, K + 1, dtype=torch.float32, device=self.device)
        nb = self.abi.mpde_maml_workspace_bytes(C.byref(net), C.byref(pdi), None, C.byref(mc), T)
        ws = self.workspace(max(nb, 1))
        self._check(self.abi.mpde_maml_adapt(C.byref(net), C.byref(pdi), C.byref(mc), T, _ptr(self._t(theta0, "theta0")),
                                             _ptr(lrs), ipts, C.byref(fpts), _ptr(prm), _ptr(thf), _ptr(losses),
                                             _ptr(ws), ws.numel(), self._stream()))
        return thf, losses

    def leap_outer(self, net, pde, bc_weight, cfg, theta0, loss0_sets, step_sets, new_sets, pde_params, T,
                   grad_scale=None, want_theta_final=False):
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
        prm = self._t(pde_params, "pde_params")
        g = torch.empty(P, dtype=torch.float32, device=self.device)
        losses = torch.empty(T, K + 1, dtype=torch.float32, device=self.device)
        thf = torch.empty(T, P, dtype=torch.float32, device=self.device) if want_theta_final else None
        nb = self.abi.mpde_leap_workspace_bytes(C.byref(net), C.byref(pd), C.byref(lc), T)
        ws = self.workspace(max(nb, 1))
        gs = (1.0 / T) if grad_scale is None else float(grad_scale)
        self._check(self.abi.mpde_leap_outer(C.byref(net), C.byref(pd), C.byref(lc), T, _ptr(self._t(theta0, "theta0")),
                                             C.byref(l0), sp, npn, _ptr(prm), gs, _ptr(g), _ptr(losses), _ptr(thf),
                                             _ptr(ws), ws.numel(), self._stream()))
        return {"meta_grad": g, "losses": losses, "theta_final": thf}
