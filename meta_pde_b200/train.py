"""Host-side mirror of the reference's jitted MAML / LEAP ``train_step`` for a flat theta.

Reference (paths relative to /root/reference):
  * MAML  ``train_step``   src/maml_pde.py:203-226  (multi_task_grad_and_losses, clip by the MODEL-grad norm,
                           outer optimizer on theta, jax.experimental.optimizers.adam on the lr tree)
  * LEAP  ``train_step``   src/leap_pde.py:175-195
  * ``get_optimizer``      src/util/trainer_util.py:593-612 ("adam": Adam(b1 .9, b2 .99), "rmsprop": Adam(0, .8),
                           "sgd": Momentum(beta 0); "ranger" is an un-pinned external package and is refused)

Everything here is launch orchestration: the arithmetic runs in libmetapde_b200.so (the fused jet kernels and
the parameter-sized kernels), the task mean across ranks is one NCCL all-reduce of the packed
``[g_theta | g_lrs]`` buffer.  torch is used for device memory, streams and
``torch.distributed`` only.  No CPU fallback: constructing a trainer without CUDA raises.
"""
import torch

from . import ops
from ._abi import PDE_NPRM, PDE_OUT_DIM, PDE_SETS, PDE_TERMS

_ADAM_BETAS = {"adam": (0.9, 0.99), "rmsprop": (0.0, 0.8)}


class MamlBatch:
    """Pre-sampled points of ONE outer step for T tasks, packed in one flat float32 buffer.

    Order of the (2K+2) point tuples: inner[0..K-1] (key k1 of scan step k, maml.py:134), outer[0..K-1] (k2),
    final (final_key, maml.py:158), aux (outer_loss_key, maml.py:192); inside a tuple the ABI set order
    (include/metapde_b200.h); each set is [T, n_s, 2].  ``pde_params`` is [T, nprm]."""

    def __init__(self, pde, T, K, n_inner, n_outer, flat, pde_params):
        self.pde, self.T, self.K = pde, int(T), int(K)
        self.n_inner, self.n_outer = list(n_inner), list(n_outer)
        self.flat, self.pde_params = flat, pde_params

    @staticmethod
    def numel(T, K, n_inner, n_outer):
        return 2 * T * ((K + 1) * sum(n_inner) + (K + 1) * sum(n_outer))

    @classmethod
    def pack(cls, pde, K, inner, outer, final, aux, pde_params, pin=False):
        """inner/outer: K lists of [T, n_s, 2] tensors; final/aux: one list each."""
        T = inner[0][0].shape[0]
        n_in = [s.shape[1] for s in inner[0]]
        n_out = [s.shape[1] for s in outer[0]]
        flat = torch.empty(cls.numel(T, K, n_in, n_out), dtype=torch.float32, pin_memory=pin)
        off = 0
        for sets in list(inner) + list(outer) + [final, aux]:
            for s in sets:
                n = s.numel()
                flat[off:off + n] = s.reshape(-1)
                off += n
        assert off == flat.numel()
        prm = pde_params.to(torch.float32).contiguous()
        if pin:
            prm = prm.pin_memory()
        return cls(pde, T, K, n_in, n_out, flat, prm)

    def to(self, device, out=None, non_blocking=True):
        """Host -> device copy (into ``out``'s buffers when given); returns a device MamlBatch."""
        if out is None:
            out = MamlBatch(self.pde, self.T, self.K, self.n_inner, self.n_outer,
                            torch.empty(self.flat.numel(), dtype=torch.float32, device=device),
                            torch.empty(self.pde_params.shape, dtype=torch.float32, device=device))
        out.flat.copy_(self.flat, non_blocking=non_blocking)
        out.pde_params.copy_(self.pde_params, non_blocking=non_blocking)
        return out

    def nbytes(self):
        return 4 * (self.flat.numel() + self.pde_params.numel())

    def views(self):
        """-> (inner [K][S], outer [K][S], final [S], aux [S]) as views of the flat buffer."""
        T, K = self.T, self.K
        off = 0
        tuples = []
        for sizes in [self.n_inner] * K + [self.n_outer] * K + [self.n_inner, self.n_outer]:
            sets = []
            for n in sizes:
                sets.append(self.flat[off:off + 2 * T * n].view(T, n, 2))
                off += 2 * T * n
            tuples.append(sets)
        return tuples[:K], tuples[K:2 * K], tuples[2 * K], tuples[2 * K + 1]


class MamlTrainer:
    """State + one outer step of MAML meta-training on this rank's share of the meta-batch.

    ``train_step(step, batch)`` reproduces src/maml_pde.py:203-226 for the tasks in ``batch`` (the local
    T_local = bsize / world_size tasks); with ``torch.distributed`` initialised the summed local gradients are
    all-reduced (NCCL over NVLink) and scaled by 1/bsize, which is the reference's mean over the vmapped tasks
    (src/nets/maml.py:231).  The clip + optimizer updates then run identically on every rank.
    """

    def __init__(self, pde, num_layers, layer_size, theta0, *, bsize, inner_steps, inner_lr, outer_lr, lr_inner_lr,
                 siren_omega=30.0, log_scale=True, bc_weight=1.0, inner_grad_clip=1e14, grad_clip=1e14,
                 outer_loss_decay=0.1, optimizer="adam", device=None, group=None, use_cuda_graph=False, kernels=None,
                 nonlinearity="sin"):
        # `kernels` is for the CPU tests, which bind the emulator build of the kernel sources; the product path is
        # ops.kernels(), which raises without CUDA / without the built library (no CPU fallback)
        self.K = kernels if kernels is not None else ops.kernels(device)
        self.device = self.K.device
        self.pde = pde
        self.net = self.K.net_desc(num_layers, layer_size, PDE_OUT_DIM[pde], siren_omega, log_scale, nonlinearity=nonlinearity)
        self.P = self.K.param_count(self.net)
        if theta0.numel() != self.P:
            raise ValueError(f"theta0 has {theta0.numel()} parameters, the network needs {self.P}")
        if optimizer not in ("adam", "rmsprop", "sgd"):
            raise Exception("unknown optimizer: ", optimizer)   # trainer_util.py:610 (ranger: external, unpinned)
        self.optimizer = optimizer
        self.bsize, self.inner_steps = int(bsize), int(inner_steps)
        self.cfg = dict(inner_steps=self.inner_steps, inner_lr=float(inner_lr), inner_grad_clip=float(inner_grad_clip),
                        softplus_lrs=True, outer_loss_decay=float(outer_loss_decay))   # MamlDef, maml_pde.py:103-110
        self.bc_weight, self.grad_clip = float(bc_weight), float(grad_clip)
        self.outer_lr, self.lr_inner_lr = float(outer_lr), float(lr_inner_lr)
        dev = self.device
        self.theta = theta0.detach().to(dev, torch.float32).clone().contiguous()
        self.inner_lrs = torch.ones(self.inner_steps, self.P, dtype=torch.float32, device=dev)   # maml_pde.py:151-154
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        self.m_theta, self.v_theta = z(self.P), z(self.P)
        self.m_lrs, self.v_lrs = z(self.inner_steps, self.P), z(self.inner_steps, self.P)
        self.t_theta = 0
        self.group = group
        self.world = torch.distributed.get_world_size(group) if torch.distributed.is_initialized() else 1
        if self.bsize % self.world:
            raise ValueError(f"bsize {self.bsize} must be a multiple of the world size {self.world}")
        self.T_local = self.bsize // self.world
        self._packed = torch.empty((self.inner_steps + 1) * self.P, dtype=torch.float32, device=dev)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self._dev_batch = None
        # The rollout workspace is PRIVATE to the trainer: a captured graph bakes its address in, and the
        # Kernels object's shared grow-only buffer may be freed by any later, larger call.
        self._ws = None
        # CUDA graph of the rollout (mpde_maml_outer: ~400 launches per step, which is what a step of the small
        # published configs -- 3x64, 1024 points -- costs).  Captured on the first step for the batch
        # shape seen there and replayed afterwards; the all-reduce and the optimizer updates (whose bias corrections
        # are host scalars that change every step) stay outside the graph.
        self.use_cuda_graph = bool(use_cuda_graph) and self.device.type == "cuda"
        self._graph = None
        self._graph_batch = None
        self._graph_out = None
        self._graph_key = None
        self.graph_launches = 0

    # -- pieces (bench.py times them separately) ------------------------------------------------------------
    def stage(self, host_batch):
        """H2D copy of a pinned host batch into the trainer's persistent device staging buffers."""
        self._dev_batch = host_batch.to(self.device, out=self._dev_batch)
        return self._dev_batch

    def _workspace(self, batch):
        nb = self.K.maml_workspace_bytes(self.net, self.pde, self.bc_weight, self.cfg, batch.n_inner, batch.n_outer, batch.T)
        if self._ws is None or self._ws.numel() < nb:
            if self._ws is not None and self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
                self._ws = None
                self._graph = None                       # a captured graph holds the old address
                torch.cuda.empty_cache()
            self._ws = torch.empty(int(nb), dtype=torch.uint8, device=self.device)
        return self._ws

    def _rollout(self, batch):
        inner, outer, final, aux = batch.views()
        P, K = self.P, self.inner_steps
        return self.K.maml_outer(self.net, self.pde, self.bc_weight, self.cfg, self.theta, self.inner_lrs, inner, outer,
                                 final, aux, batch.pde_params, batch.T, grad_scale=1.0 / self.bsize,
                                 out_g_theta=self._packed[:P], out_g_lrs=self._packed[P:].view(K, P),
                                 workspace=self._workspace(batch))

    def _rollout_graphed(self, batch):
        key = (batch.T, tuple(batch.n_inner), tuple(batch.n_outer), self._workspace(batch).data_ptr())
        if self._graph is None or self._graph_key != key:
            # static input buffers; one eager call first (workspace allocation, one-time function attributes), then
            # capture on a side stream as torch.cuda.graph requires
            self._graph_batch = MamlBatch(batch.pde, batch.T, batch.K, batch.n_inner, batch.n_outer,
                                          torch.empty_like(batch.flat), torch.empty_like(batch.pde_params))
            self._graph_batch.flat.copy_(batch.flat)
            self._graph_batch.pde_params.copy_(batch.pde_params)
            self._rollout(self._graph_batch)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            n0 = self.K.launch_count()
            with torch.cuda.graph(g):
                self._graph_out = self._rollout(self._graph_batch)
            self.graph_launches = self.K.launch_count() - n0     # kernel nodes replayed per step
            self._graph, self._graph_key = g, key
        if batch is not self._graph_batch:
            self._graph_batch.flat.copy_(batch.flat, non_blocking=True)
            self._graph_batch.pde_params.copy_(batch.pde_params, non_blocking=True)
        self._graph.replay()
        return self._graph_out

    def meta_grad(self, batch):
        """multi_task_grad_and_losses (maml.py:210-233) over the local tasks; returns the ABI result dict with
        g_theta / g_lrs holding (1/bsize) * SUM over local tasks (all-reduced when distributed)."""
        out = self._rollout_graphed(batch) if self.use_cuda_graph else self._rollout(batch)
        if self.world > 1:
            torch.distributed.all_reduce(self._packed, group=self.group)
        return out

    def apply(self, step, out=None):
        """Clip by the model-grad norm (quirk: the lr-tree grads are rescaled by it too) + both optimizer updates:
        ONE launch of the fused outer-step kernel over the packed (all-reduced) buffer.  Returns sum(g_theta^2)."""
        self.t_theta += 1
        b1, b2 = _ADAM_BETAS.get(self.optimizer, (0.9, 0.99))
        # jax.experimental.optimizers.adam(lr_inner_lr, b1=.9, b2=.99) at step index `step` (maml_pde.py:150,225)
        return self.K.outer_step(self.theta, self._packed, self.m_theta, self.v_theta, self.t_theta, self.outer_lr, b1, b2,
                                 1e-8, self.grad_clip, self.optimizer, lrs=self.inner_lrs.view(-1), m_lrs=self.m_lrs.view(-1),
                                 v_lrs=self.v_lrs.view(-1), t_lrs=int(step) + 1, lr_lrs=self.lr_inner_lr, out=self._sumsq)

    def train_step(self, step, batch):
        """One outer step on a DEVICE batch.  Returns (losses [T_local, K+1], meta_losses [T_local],
        aux_terms [T_local, n_terms], meta_grad_norm [1]) as device tensors (no host sync)."""
        out = self.meta_grad(batch)
        sumsq = self.apply(step)
        return out["losses"], out["meta_losses"], out["aux_terms"], torch.sqrt(sumsq)

    def train_step_host(self, step, host_batch):
        """The end-to-end call: pinned host batch in, host results out (H2D + step + D2H)."""
        losses, meta, aux, gnorm = self.train_step(step, self.stage(host_batch))
        res = torch.cat([losses.reshape(-1), meta.reshape(-1), aux.reshape(-1), gnorm.reshape(-1)]).cpu()
        T, K = losses.shape[0], self.inner_steps
        nt = len(PDE_TERMS[self.pde])
        return (res[:T * (K + 1)].view(T, K + 1), res[T * (K + 1):T * (K + 2)],
                res[T * (K + 2):T * (K + 2 + nt)].view(T, nt), res[-1])


def synthetic_maml_batch(pde, T, K, n_inner, n_outer, seed, pin=False, task_offset=0, T_global=None):
    """Seeded synthetic MamlBatch: tasks [task_offset, task_offset + T) of the GLOBAL batch of
    ``T_global`` tasks drawn from ``seed`` -- every rank count sees the same task set."""
    from . import synthetic

    Tg = int(T_global) if T_global is not None else int(task_offset) + int(T)
    si, so = synthetic.default_set_sizes(pde, n_inner), synthetic.default_set_sizes(pde, n_outer)
    base = int(seed) * 100003
    sl = slice(int(task_offset), int(task_offset) + int(T))
    cut = lambda sets: [s[sl].contiguous() for s in sets]
    inner, outer = [], []
    _, prm = synthetic.make_tasks(pde, Tg, [1] * len(PDE_SETS[pde]), seed=base)
    for k in range(K):
        inner.append(cut(synthetic.make_tasks(pde, Tg, si, seed=base + 10 * k + 1)[0]))
        outer.append(cut(synthetic.make_tasks(pde, Tg, so, seed=base + 10 * k + 2)[0]))
    final = cut(synthetic.make_tasks(pde, Tg, si, seed=base + 1001)[0])
    aux = cut(synthetic.make_tasks(pde, Tg, so, seed=base + 1002)[0])
    assert prm.shape[1] == PDE_NPRM[pde]
    return MamlBatch.pack(pde, K, inner, outer, final, aux, prm[sl].contiguous(), pin=pin)


class PointBatch:
    """Pre-sampled point tuples of one step for T tasks in ONE flat float32 buffer (one H2D copy): ``groups`` is a list
    of point tuples, each a list of [T, n_s, 2] tensors in the ABI set order."""

    def __init__(self, T, sizes, n_groups, flat, pde_params):
        self.T, self.sizes, self.n_groups, self.flat, self.pde_params = int(T), list(sizes), int(n_groups), flat, pde_params

    @classmethod
    def pack(cls, groups, pde_params, pin=False):
        T = groups[0][0].shape[0]
        sizes = [s.shape[1] for s in groups[0]]
        flat = torch.empty(2 * T * sum(sizes) * len(groups), dtype=torch.float32, pin_memory=pin)
        off = 0
        for sets in groups:
            assert [s.shape[1] for s in sets] == sizes
            for s in sets:
                flat[off:off + s.numel()] = s.reshape(-1)
                off += s.numel()
        prm = pde_params.to(torch.float32).contiguous()
        return cls(T, sizes, len(groups), flat, prm.pin_memory() if pin else prm)

    def to(self, device, out=None, non_blocking=True):
        if out is None:
            out = PointBatch(self.T, self.sizes, self.n_groups, torch.empty(self.flat.numel(), dtype=torch.float32, device=device),
                             torch.empty(self.pde_params.shape, dtype=torch.float32, device=device))
        out.flat.copy_(self.flat, non_blocking=non_blocking)
        out.pde_params.copy_(self.pde_params, non_blocking=non_blocking)
        return out

    def nbytes(self):
        return 4 * (self.flat.numel() + self.pde_params.numel())

    def views(self):
        off, groups = 0, []
        for _ in range(self.n_groups):
            sets = []
            for n in self.sizes:
                sets.append(self.flat[off:off + 2 * self.T * n].view(self.T, n, 2))
                off += 2 * self.T * n
            groups.append(sets)
        return groups


class _OuterOpt:
    """theta + the outer optimizer state of get_optimizer (src/util/trainer_util.py:593-612) + the fused update."""

    def _init_opt(self, theta0, optimizer, outer_lr, grad_clip):
        if optimizer not in ("adam", "rmsprop", "sgd"):
            raise Exception("unknown optimizer: ", optimizer)
        dev = self.device
        self.optimizer, self.outer_lr, self.grad_clip = optimizer, float(outer_lr), float(grad_clip)
        self.theta = theta0.detach().to(dev, torch.float32).clone().contiguous()
        self.m_theta, self.v_theta = torch.zeros_like(self.theta), torch.zeros_like(self.theta)
        self.t_theta = 0
        self._g = torch.empty(self.P, dtype=torch.float32, device=dev)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self._dev_batch = None

    def apply(self):
        self.t_theta += 1
        b1, b2 = _ADAM_BETAS.get(self.optimizer, (0.9, 0.99))
        return self.K.outer_step(self.theta, self._g, self.m_theta, self.v_theta, self.t_theta, self.outer_lr, b1, b2, 1e-8,
                                 self.grad_clip, self.optimizer, out=self._sumsq)

    def stage(self, host_batch):
        self._dev_batch = host_batch.to(self.device, out=self._dev_batch)
        return self._dev_batch


class LeapTrainer(_OuterOpt):
    """One outer step of LEAP meta-training (src/leap_pde.py:175-195) on this rank's share of the meta-batch:
    mpde_leap_outer (first-order rollout + increments, src/nets/leap.py:49-228), NCCL all-reduce of the task sum,
    fused clip + optimizer step.  Batch = PointBatch with groups [loss0, step_0..K-1, new_0..K-1]."""

    def __init__(self, pde, num_layers, layer_size, theta0, *, bsize, inner_steps, inner_lr, outer_lr, siren_omega=30.0,
                 log_scale=True, bc_weight=1.0, inner_grad_clip=1e14, grad_clip=1e14, optimizer="adam", inner_opt="adam",
                 device=None, group=None, kernels=None, nonlinearity="sin"):
        self.K = kernels if kernels is not None else ops.kernels(device)
        self.device, self.pde = self.K.device, pde
        self.net = self.K.net_desc(num_layers, layer_size, PDE_OUT_DIM[pde], siren_omega, log_scale, nonlinearity=nonlinearity)
        self.P = self.K.param_count(self.net)
        self.bsize, self.inner_steps, self.bc_weight = int(bsize), int(inner_steps), float(bc_weight)
        self.cfg = dict(inner_steps=self.inner_steps, inner_lr=float(inner_lr), inner_grad_clip=float(inner_grad_clip),
                        norm=True, loss_in_distance=True, stabilize=True, inner_opt=inner_opt, beta1=0.9, beta2=0.99,
                        eps=1e-8)                                                     # LeapDef, leap_pde.py:87-97
        self._init_opt(theta0, optimizer, outer_lr, grad_clip)
        self.group = group
        self.world = torch.distributed.get_world_size(group) if torch.distributed.is_initialized() else 1
        if self.bsize % self.world:
            raise ValueError(f"bsize {self.bsize} must be a multiple of the world size {self.world}")

    def train_step(self, batch):
        g = batch.views()
        K = self.inner_steps
        out = self.K.leap_outer(self.net, self.pde, self.bc_weight, self.cfg, self.theta, g[0], g[1:1 + K], g[1 + K:1 + 2 * K],
                                batch.pde_params, batch.T, grad_scale=1.0 / self.bsize, out_meta_grad=self._g)
        if self.world > 1:
            torch.distributed.all_reduce(self._g, group=self.group)
        sumsq = self.apply()
        return out["losses"], torch.sqrt(sumsq)

    def train_step_host(self, host_batch):
        losses, gnorm = self.train_step(self.stage(host_batch))
        res = torch.cat([losses.reshape(-1), gnorm.reshape(-1)]).cpu()
        return res[:-1].view(losses.shape), res[-1]


class FinetuneTrainer(_OuterOpt):
    """The fine-tune step of the README command (src/nn_pde_maml.py:110-124,140-171): gradient of the loss SUMMED over
    the bsize tasks of the batch, clip, Adam.  Batch = PointBatch with one group.  T = bsize tasks do not shard
    (bsize = 1 in the published command): replicas only."""

    def __init__(self, pde, num_layers, layer_size, theta0, *, bsize, outer_lr, siren_omega=30.0, log_scale=True,
                 bc_weight=1.0, grad_clip=1e14, optimizer="adam", device=None, kernels=None, nonlinearity="sin"):
        self.K = kernels if kernels is not None else ops.kernels(device)
        self.device, self.pde = self.K.device, pde
        self.net = self.K.net_desc(num_layers, layer_size, PDE_OUT_DIM[pde], siren_omega, log_scale, nonlinearity=nonlinearity)
        self.P = self.K.param_count(self.net)
        self.bsize, self.bc_weight = int(bsize), float(bc_weight)
        self._init_opt(theta0, optimizer, outer_lr, grad_clip)

    def train_step(self, batch):
        (sets,) = batch.views()
        loss, terms, grad = self.K.loss_grad(self.net, self.pde, self.bc_weight, self.theta, sets, batch.pde_params, batch.T)
        self.K.task_sum(grad, 1.0, out=self._g)
        sumsq = self.apply()
        return loss, terms, torch.sqrt(sumsq)

    def train_step_host(self, host_batch):
        loss, terms, gnorm = self.train_step(self.stage(host_batch))
        res = torch.cat([loss.reshape(-1), terms.reshape(-1), gnorm.reshape(-1)]).cpu()
        T = loss.shape[0]
        return res[:T].sum(), res[T:-1].view(terms.shape).mean(dim=0), res[-1]


def synthetic_point_batch(pde, T, n_groups, n, seed, pin=False, task_offset=0, T_global=None):
    """Seeded synthetic PointBatch: ``n_groups`` fresh point tuples of size ``n`` for tasks [task_offset, +T) of a
    global batch of T_global tasks (same task set for every rank count)."""
    from . import synthetic

    Tg = int(T_global) if T_global is not None else int(task_offset) + int(T)
    sizes = synthetic.default_set_sizes(pde, n)
    base = int(seed) * 100003 + 17
    sl = slice(int(task_offset), int(task_offset) + int(T))
    _, prm = synthetic.make_tasks(pde, Tg, [1] * len(PDE_SETS[pde]), seed=base)
    groups = [[s[sl].contiguous() for s in synthetic.make_tasks(pde, Tg, sizes, seed=base + 1 + j)[0]] for j in range(n_groups)]
    return PointBatch.pack(groups, prm[sl].contiguous(), pin=pin)
