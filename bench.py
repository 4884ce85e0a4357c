#!/usr/bin/env python
"""bench.py -- the hot path of meta-pde on B200: MAML outer steps/s on BASELINE.json's scaled Poisson config (C5,
the default), and the published small configs C1-C4 (README fine-tune step, LEAP, Burgers MAML, hyperelasticity MAML).

    python bench.py --gpus N --steps K --warmup W          (N > 1: launched by torchrun, one rank per GPU)
    python bench.py --workload c1_poisson_finetune_3x64_1k ...
    python bench.py --impl reference ...                   (the CPU arm: the oracle port on the host cores)

One "step" = one full outer step over the whole meta-batch:
  maml      src/maml_pde.py:203-226: per task K inner SGD steps (loss+grad), K outer loss+grads, K Hessian-vector
            products, 2 loss-only evals, task mean (NCCL all-reduce of [g_theta | g_lrs] when N > 1),
            clip + Adam(theta) + Adam(alpha) in one fused launch
  leap      src/leap_pde.py:175-195: per task K x (loss+grad, Adam inner step, loss) + increments, task mean, clip + Adam
  finetune  src/nn_pde_maml.py:140-171 (the README command's steady-state step): loss+grad of the batch, clip + Adam
Nothing is skipped or cached inside the timed region.  The meta-batch is sharded over the ranks (T/N tasks per GPU:
"strong" scaling, the config fixes the global meta-batch); the T = 1 fine-tune step does not shard (replicas only).

Printed JSON line (rank 0): `value` = the metric with this step's points already in HBM (per-kernel event timing OFF),
`e2e` = the same step through the trainer's train_step_host with pinned HOST buffers (H2D of the step's points and D2H
of its losses inside the timed region), `roofline` = the dominant kernel timed live with CUDA events on its launch
stream in a separate short pass, `cpu_baseline` = the oracle timed on the host cores (rank 0, N = 1 only),
`checksum` = float64 (sum, L2) of theta and of the last packed gradient: identical seeds at every N, so the lines of
different N are comparable (1-GPU == N-GPU up to the all-reduce's summation order).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# ---------------------------------------------------------------------------------------------- workloads
WORKLOADS = {
    # BASELINE.json configs[4]: "Scaled Poisson MAML sweep: meta-batch 64 tasks x 16k points, SIREN 5x256"
    "c5_poisson_maml_5x256_T64_16k": dict(kind="maml", pde="poisson", L=5, W=256, T=64, n=16384, K=5),
    # configs[0]: the README command (src/nn_pde_maml, /root/reference/README.md:22): 3x64, 1024 + 1024 points, one task;
    # the only published figure of the path: 0.01599 s per step (src/notebooks/resultAnalysis.ipynb:443)
    "c1_poisson_finetune_3x64_1k": dict(kind="finetune", pde="poisson", L=3, W=64, T=1, n=1024, K=0),
    # configs[1]: Poisson LEAP 3x64, 1024 points; bsize 16 and inner_steps 10 are the driver defaults (src/leap_pde.py:42,48)
    "c2_poisson_leap_3x64_T16_1k": dict(kind="leap", pde="poisson", L=3, W=64, T=16, n=1024, K=10),
    # configs[2] / [3]: Burgers / hyperelasticity MAML 3x64, T = 8, K = 5 (pipeline/pipleline_maml_meta.sh:5-6)
    "c3_burgers_maml_3x64_T8_1k": dict(kind="maml", pde="td_burgers", L=3, W=64, T=8, n=1024, K=5),
    "c4_hyper_maml_3x64_T8_1k": dict(kind="maml", pde="hyper_elasticity", L=3, W=64, T=8, n=1024, K=5),
    # the Poisson MAML pipeline shape (pipeline/pipleline_maml_meta.sh:4) and small shapes for smoke runs of this script
    "dev_poisson_maml_3x64_T16_1k": dict(kind="maml", pde="poisson", L=3, W=64, T=16, n=1024, K=5),
    "dev_poisson_maml_5x256_T8_2k": dict(kind="maml", pde="poisson", L=5, W=256, T=8, n=2048, K=5),
}
DEFAULT_WORKLOAD = "c5_poisson_maml_5x256_T64_16k"


def initial_theta(wl, hy, od):
    """SIREN init (field.py:23-44), or the variance-scaling init of the --nosiren network (field.py:171-176)."""
    from meta_pde_b200 import synthetic
    if wl.get("act", "sin") == "swish":
        return synthetic.swish_theta([wl["W"]] * wl["L"], od, True, seed=0)
    return synthetic.siren_theta([wl["W"]] * wl["L"], od, hy["omega"], hy["omega"], True, seed=0)
# pipeline/pipleline_maml_meta.sh:4 (Poisson): inner_lr 1e-4, outer_lr 1e-5, clips 100, bc_weight 1; lr_inner_lr and
# outer_loss_decay are the flag defaults (src/maml_pde.py:52-58).  Hyperelasticity uses inner_lr 1e-5 (:6).
HYPER = dict(inner_lr=1e-4, outer_lr=1e-5, lr_inner_lr=0.5, inner_grad_clip=100.0, grad_clip=100.0,
             outer_loss_decay=0.1, bc_weight=1.0, omega=30.0)
METRIC = {"maml": ("maml_outer_steps_per_sec", "steps/s"), "leap": ("leap_outer_steps_per_sec", "steps/s"),
          "finetune": ("residual_grad_evals_per_sec", "domain-point loss+grad evals/s")}


def hyper_for(wl):
    h = dict(HYPER)
    if wl["pde"] == "hyper_elasticity":
        h["inner_lr"] = 1e-5
    if wl["kind"] == "leap":
        h["inner_lr"], h["outer_lr"] = 2.5e-5, 5.0e-5      # pipeline/pipeline_leap_meta.sh:5
    return h


def out_dim(pde):
    return 2 if pde == "hyper_elasticity" else 1


def weight_count(W, L, out=1):
    return 2 * W + (L - 1) * W * W + W * out


def jet_rows(pde, sizes):
    C = 3 if pde == "hyper_elasticity" else 4
    return C * sizes[-1] + sum(sizes[:-1])


def step_flops(wl, sizes):
    """Algorithmic flops of one step over all T tasks: MAML (15K+2) * 2 * P_w * R per task,
    LEAP (8K+2) * P_w * R, loss+grad 6 * P_w * R."""
    per = {"maml": (15 * wl["K"] + 2) * 2.0, "leap": 8.0 * wl["K"] + 2.0, "finetune": 6.0}[wl["kind"]]
    return per * weight_count(wl["W"], wl["L"], out_dim(wl["pde"])) * jet_rows(wl["pde"], sizes) * wl["T"]


def metric_value(wl, sizes, ms_step):
    """steps/s, or for the fine-tune step the domain points whose residual + gradient one step evaluates, per second"""
    if wl["kind"] == "finetune":
        return wl["T"] * sizes[-1] / (ms_step * 1e-3)
    return 1e3 / ms_step


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.th = index, [], None, None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.th.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        # "under load": ignore idle samples (below half of the top sample)
        load = [x for x in sm if x >= 0.5 * sm[-1]] if sm else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------- CPU baseline
def _oracle_step_fn(wl, sample_tasks, n):
    """-> one_step(seed): seconds of ONE step of the workload's kind on `sample_tasks` tasks x n points through the
    oracle (oracle/: torch.func restatement of the reference's JAX math, float32 like the reference)."""
    import torch
    from meta_pde_b200 import synthetic
    from oracle import meta as ometa
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_params, oracle_points

    pde, L, W, K, kind = wl["pde"], wl["L"], wl["W"], wl["K"], wl["kind"]
    hy = hyper_for(wl)
    od = out_dim(pde)
    lf = ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, hy["omega"], True, od, wl.get("act", "sin")), hy["bc_weight"])
    theta = initial_theta(wl, hy, od)
    sizes = synthetic.default_set_sizes(pde, n)
    f32 = torch.float32
    mk = lambda s: synthetic.make_tasks(pde, sample_tasks, sizes, seed=s)[0]
    pts = lambda sets, t: oracle_points(pde, sets, t, f32)

    def one_step(seed):
        prm = synthetic.make_tasks(pde, sample_tasks, sizes, seed=seed)[1]
        prms = [oracle_params(pde, prm, t, f32) for t in range(sample_tasks)]
        if kind == "maml":
            cfg = ometa.MamlCfg(K, hy["inner_lr"], hy["inner_grad_clip"], True, hy["outer_loss_decay"])
            lrs = torch.ones(K, theta.numel())
            inner = [mk(seed + 10 * k + 1) for k in range(K)]
            outer = [mk(seed + 10 * k + 2) for k in range(K)]
            final, aux = mk(seed + 1001), mk(seed + 1002)
            tasks = [ometa.MamlTask(prms[t], [pts(s, t) for s in inner], [pts(s, t) for s in outer], pts(final, t), pts(aux, t))
                     for t in range(sample_tasks)]
            t0 = time.perf_counter()
            ometa.maml_multi_task_grad_and_losses(cfg, lf, theta, tasks, lrs)
        elif kind == "leap":
            cfg = ometa.LeapCfg(K, hy["inner_lr"], hy["inner_grad_clip"], True, True, True, "adam")
            l0 = mk(seed + 1)
            st = [mk(seed + 20 + k) for k in range(K)]
            nw = [mk(seed + 60 + k) for k in range(K)]
            tasks = [ometa.LeapTask(prms[t], pts(l0, t), [pts(s, t) for s in st], [pts(s, t) for s in nw]) for t in range(sample_tasks)]
            t0 = time.perf_counter()
            ometa.leap_multi_task_grad_and_losses(cfg, lf, theta, tasks)
        else:
            sets = mk(seed + 1)
            batch = [(pts(sets, t), prms[t]) for t in range(sample_tasks)]
            opt = ometa.FlaxAdam(hy["outer_lr"], 0.9, 0.99)
            t0 = time.perf_counter()
            ometa.finetune_train_step(lf, theta, batch, opt, opt.init(theta), hy["grad_clip"])
        return time.perf_counter() - t0

    return one_step, sizes


def _time_oracle(wl, sample_tasks, n, threads, steps, warmup, budget_s):
    import torch
    torch.set_num_threads(threads)
    one_step, sizes = _oracle_step_fn(wl, sample_tasks, n)
    for i in range(warmup):
        one_step(100 + i)
    times = []
    for i in range(steps):
        times.append(one_step(200 + i))
        if sum(times) > budget_s:
            break
    times.sort()
    return times[len(times) // 2], len(times), sizes


def cpu_baseline(wl, steps=4, warmup=1, budget_s=30.0, one_thread=True):
    """The oracle timed on the host cores.  C1-C4 run at FULL size (all tasks, all points: no extrapolation).  C5 (64 tasks
    x 16384 + 16384 points of the 5x256 net) is sampled as ONE task at the full point count when the host has the memory
    for its second-order graph (else 4096 points), and only the task factor (x64; tasks are independent) is extrapolated.
    Both an all-cores figure and a 1-thread figure (the reference pinned its CPU baselines to one core with
    `taskset -c 0`, pipeline/pipeline_baseline.sh:3-9) are reported; `value` is the all-cores one."""
    cores = os.cpu_count() or 1
    full = wl["T"] * wl["n"] <= 16 * 1024
    if full:
        st, n = wl["T"], wl["n"]
    else:
        st, n = 1, wl["n"]
        try:
            import psutil
            if psutil.virtual_memory().available < 120e9 and n > 4096:
                n = 4096
        except Exception:
            n = min(n, 4096)
    t_all, reps, sizes = _time_oracle(wl, st, n, cores, steps, warmup, budget_s)
    scale = (wl["T"] / st) * (wl["n"] / n)
    val = lambda t: metric_value(wl, _sizes_full(wl), t * scale * 1e3)
    out = {"value": val(t_all), "unit": METRIC[wl["kind"]][1], "cores": cores, "kind": "port",
           "sample": (f"{st} task(s) x {n} points per set of the {wl['L']}x{wl['W']} net, one {wl['kind']} step"
                      + (f" (K={wl['K']}, full second-order outer gradient)" if wl["kind"] == "maml" else "")
                      + f", float32 torch.func oracle; median of {reps} repetition(s) after {warmup} warm-up, {t_all:.3f} s each; "
                      + ("full size, not extrapolated" if scale == 1 else f"scaled x{scale:g} (independent tasks x points, linear)")),
           "sample_seconds": t_all, "steps_timed": reps, "extrapolation_factor": scale}
    if one_thread:
        # 1 thread: a smaller sample where the full one would take minutes (stated; linear in tasks x points)
        st1, n1 = (st, n) if full and wl["kind"] != "maml" else (1, min(n, 1024))
        t1, reps1, _ = _time_oracle(wl, st1, n1, 1, 1, 0, budget_s)
        sc1 = (wl["T"] / st1) * (wl["n"] / n1)
        out["one_thread"] = {"value": metric_value(wl, _sizes_full(wl), t1 * sc1 * 1e3), "cores": 1,
                             "sample": f"{st1} task(s) x {n1} points, 1 repetition, {t1:.3f} s, scaled x{sc1:g}"}
    return out


def _sizes_full(wl):
    from meta_pde_b200 import synthetic
    return synthetic.default_set_sizes(wl["pde"], wl["n"])


def run_reference(args, wl, name):
    """--impl reference: the reference's CPU implementation of the path.  JAX/Flax are not installable here
    (DESIGN.md), so this is the oracle port, all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_baseline(wl, steps=max(1, args.steps), warmup=min(args.warmup, 1), budget_s=120.0, one_thread=False)
    metric, unit = METRIC[wl["kind"]]
    ms_step = cb["sample_seconds"] * cb["extrapolation_factor"] * 1e3
    line = {"impl": "reference", "metric": metric, "value": cb["value"], "unit": unit,
            "n_gpus": args.gpus, "steps": cb["steps_timed"], "warmup": min(args.warmup, 1),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(wl, name, args.gpus),
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def bench_config(wl, name, n_gpus):
    shards = wl["T"] % n_gpus == 0 and wl["kind"] != "finetune"
    return {"workload": name, "kind": wl["kind"], "pde": wl["pde"], "siren": (f"{wl['L']}x{wl['W']} omega=30" if wl.get("act", "sin") == "sin" else f"--nosiren: {wl['L']}x{wl['W']} swish"),
            "meta_batch_tasks": wl["T"], "tasks_per_gpu": wl["T"] // n_gpus if shards else wl["T"],
            "points_per_set": wl["n"], "inner_steps": wl["K"], "second_order": wl["kind"] == "maml",
            "parallelism": f"task-sharded dp{n_gpus}" if shards else f"{n_gpus} independent replica(s)",
            "cuda_graph": bool(getattr(bench_config, "graph", False)),
            "l2_policy": ("two alternating pre-staged batches of fresh point sets; per-step working set (activation spills, "
                          "weight images) >> 126 MB L2 at C5; small configs: L2-resident by nature (the whole step's data is < L2), "
                          "an L2 flush (256 MB write) between timed steps is added with --flush-l2")}


# ------------------------------------------------------------------------------------------------ GPU arm
def build_job(wl, dev, rank, world, args):
    """-> dict(trainer, host (2 pinned host batches), dev_step(i, step_no), host_step(i, step_no), api)"""
    import torch
    from meta_pde_b200 import synthetic, train

    pde, L, W, T, n, K, kind = (wl[k] for k in ("pde", "L", "W", "T", "n", "K", "kind"))
    hy = hyper_for(wl)
    shards = kind != "finetune"
    T_local = T // world if shards else T
    off = rank * T_local if shards else 0
    theta0 = initial_theta(wl, hy, out_dim(pde))
    act = wl.get("act", "sin")
    if kind == "maml":
        tr = train.MamlTrainer(pde, L, W, theta0, bsize=T, inner_steps=K, inner_lr=hy["inner_lr"], outer_lr=hy["outer_lr"],
                               lr_inner_lr=hy["lr_inner_lr"], siren_omega=hy["omega"], bc_weight=hy["bc_weight"],
                               inner_grad_clip=hy["inner_grad_clip"], grad_clip=hy["grad_clip"],
                               outer_loss_decay=hy["outer_loss_decay"], device=dev, use_cuda_graph=args.graph,
                               nonlinearity=act)
        host = [train.synthetic_maml_batch(pde, T_local, K, n, n, seed=s, pin=True, task_offset=off, T_global=T) for s in (1, 2)]
        dev_step = lambda b, s: tr.train_step(s, b)
        host_step = lambda b, s: tr.train_step_host(s, b)
        api = "meta_pde_b200.train.MamlTrainer.train_step_host"
        packed = lambda: tr._packed
    elif kind == "leap":
        tr = train.LeapTrainer(pde, L, W, theta0, bsize=T, inner_steps=K, inner_lr=hy["inner_lr"], outer_lr=hy["outer_lr"],
                               siren_omega=hy["omega"], bc_weight=hy["bc_weight"], inner_grad_clip=hy["inner_grad_clip"],
                               grad_clip=hy["grad_clip"], device=dev, nonlinearity=act)
        host = [train.synthetic_point_batch(pde, T_local, 1 + 2 * K, n, seed=s, pin=True, task_offset=off, T_global=T) for s in (1, 2)]
        dev_step = lambda b, s: tr.train_step(b)
        host_step = lambda b, s: tr.train_step_host(b)
        api = "meta_pde_b200.train.LeapTrainer.train_step_host"
        packed = lambda: tr._g
    else:
        tr = train.FinetuneTrainer(pde, L, W, theta0, bsize=T, outer_lr=hy["outer_lr"], siren_omega=hy["omega"],
                                   bc_weight=hy["bc_weight"], grad_clip=hy["grad_clip"], device=dev, nonlinearity=act)
        host = [train.synthetic_point_batch(pde, T_local, 1, n, seed=s, pin=True) for s in (1, 2)]
        dev_step = lambda b, s: tr.train_step(b)
        host_step = lambda b, s: tr.train_step_host(b)
        api = "meta_pde_b200.train.FinetuneTrainer.train_step_host"
        packed = lambda: tr._g
    return dict(trainer=tr, host=host, dev_step=dev_step, host_step=host_step, api=api, packed=packed, T_local=T_local,
                shards=shards)


def run_ours(args, wl, name):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this package has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # stdout carries exactly one JSON line: library chatter (e.g. NCCL's version banner) goes to stderr
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from meta_pde_b200 import synthetic
    from meta_pde_b200._abi import MODE_GRAD, MODE_HVP, MODE_LOSS

    pde, L, W, T, n, K, kind = (wl[k] for k in ("pde", "L", "W", "T", "n", "K", "kind"))
    if kind != "finetune" and T % world:
        raise SystemExit(f"meta-batch {T} does not divide over {world} ranks")
    job = build_job(wl, dev, rank, world, args)
    tr, host, T_local = job["trainer"], job["host"], job["T_local"]
    Kn = tr.K
    sizes = synthetic.default_set_sizes(pde, n)
    devb = [hb.to(dev) for hb in host]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if args.flush_l2 else None
    torch.cuda.synchronize()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        """EXACTLY `steps` steps between two events on the launch stream; with --flush-l2 each step is bracketed by its
        own event pair and a 256 MB write runs between the pairs (outside them)."""
        sync_all()
        if flush is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            sync_all()
            total = e0.elapsed_time(e1)
        else:
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for i in range(steps):
                flush.fill_(i & 0xff)
                ev[i][0].record()
                fn(i)
                ev[i][1].record()
            sync_all()
            total = sum(a.elapsed_time(b) for a, b in ev)
        ms = torch.tensor([total], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    step_no = [0]

    def dev_step(i):
        job["dev_step"](devb[i % 2], step_no[0])
        step_no[0] += 1

    last = {}

    def host_step(i):
        last["out"] = job["host_step"](host[i % 2], step_no[0])
        step_no[0] += 1

    # ---- warm-up, then EXACTLY `steps` timed steps with the points resident in HBM; per-kernel event timing OFF -------
    Kn.prof_enable(False)
    for i in range(args.warmup):
        dev_step(i)
    launches0 = Kn.launch_count()
    with ClockSampler(local) as clk:
        ms_total = timed(dev_step, args.steps)
    graph_launches = getattr(tr, "graph_launches", 0)
    launches = Kn.launch_count() - launches0 + (graph_launches * args.steps if args.graph else 0)
    ms_step = ms_total / args.steps
    # ---- end to end: pinned host points -> H2D -> step -> D2H of the losses -------------------------------
    host_step(0)
    ms_e2e = timed(host_step, args.steps) / args.steps
    out = last["out"]
    finite = bool(torch.isfinite(out[0]).all())
    # ---- checksum (comparable across N: same seeds, same number of steps) --------------------------------------
    th64, g64 = tr.theta.double(), job["packed"]().double()
    checksum = {"theta_sum": float(th64.sum()), "theta_l2": float(th64.norm()), "grad_sum": float(g64.sum()),
                "grad_l2": float(g64.norm()), "after_steps": step_no[0]}
    # ---- per-kernel timing pass (separate from the pass that produced `value`) -----------------------------
    prof_steps = min(args.steps, 3)
    prof, wgrad_ms, wgrad_n, ms_prof = {}, 0.0, 0, None
    if not args.graph:
        Kn.prof_enable(True)
        ms_prof = timed(dev_step, prof_steps)
        prof = {m: Kn.prof_read(m) for m in (MODE_LOSS, MODE_GRAD, MODE_HVP)}
        wgrad_ms, wgrad_n = Kn.prof_read(3)
        Kn.prof_enable(False)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel -----------------------------------------------------------------------
    Pw, R = weight_count(W, L, out_dim(pde)), jet_rows(pde, sizes)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    fp32_peak, fp32_src = 71.6, "profiles/r01_microbench_pipes.jsonl (FFMA microbench on this pool's B200, 296 CTAs)"
    tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    tc = W in (128, 256) and os.environ.get("MPDE_DISABLE_TC", "0") != "1" and wl.get("act", "sin") == "sin"
    dom_mode = MODE_HVP if kind == "maml" else MODE_GRAD
    roofline = None
    if prof:
        d_ms, d_n = prof[dom_mode]
        share = {k: (v[0] / ms_prof) for k, v in (("hvp", prof[MODE_HVP]), ("grad", prof[MODE_GRAD]), ("loss", prof[MODE_LOSS]))}
        share["tc_wgrad"] = wgrad_ms / ms_prof
        per_mode = {"hvp": list(prof[MODE_HVP]), "grad": list(prof[MODE_GRAD]), "loss": list(prof[MODE_LOSS]), "tc_wgrad": [wgrad_ms, wgrad_n]}
        # dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed `ncu --set full` capture of the same
        # kernel at this workload (profiles/ncu_traffic.json, written by tools/gpu_ncu_full.sh); null for other workloads
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            if tj.get("workload") == name and world == 1:
                key = ("tc_hvp_kernel" if tc else "pde_kernel_hvp") if dom_mode == MODE_HVP else ("tc_pde_kernel_grad" if tc else "pde_kernel_grad")
                traffic = tj["kernels"].get(key, {}).get("traffic_bytes")
        except (OSError, ValueError, KeyError):
            pass
        avg_ms = d_ms / max(d_n, 1)
        if tc:
            # main tensor-core kernel: 6 (HVP) / 2 (GRAD) GEMM-equivalents of 2 P_w flop per jet row as fp16 hi/lo 3-term
            # splits on tcgen05 (3 tensor-core flops per algorithmic flop); the weight-gradient GEMMs are tc_wgrad_kernel's
            k_flops = (12.0 if dom_mode == MODE_HVP else 4.0) * Pw * R * T_local
            achieved = k_flops / (avg_ms * 1e-3) / 1e12 if d_n else None
            roofline = {
                "kernel": ("tc_hvp_kernel<W,PDE>" if dom_mode == MODE_HVP else "tc_pde_kernel<W,PDE,GRAD>")
                          + " (tcgen05 kind::f16, fp16 hi/lo 3-term split, TMEM accumulators)",
                "bound": "tensor", "achieved": achieved, "peak": tensor_peak, "unit": "TFLOP/s",
                "frac": (achieved / tensor_peak) if achieved else None,
                "peak_source": "of measured: MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)",
                "achieved_issued_tflops": (3.0 * achieved) if achieved else None,
                "frac_issued": (3.0 * achieved / tensor_peak) if achieved else None,
                "note": "achieved = ALGORITHMIC fp32 flops; each is 3 fp16 tensor-core flops (a_hi*b_hi + a_lo*b_hi + a_hi*b_lo), "
                        "so frac_issued = 3*frac is the tensor-pipe occupancy and 1/3 is the ceiling of frac at fp32 accuracy",
                "traffic": traffic, "flops_per_launch": k_flops, "avg_launch_ms": avg_ms, "launches_timed": d_n,
                "kernel_share_of_step": share, "per_mode_ms": per_mode,
            }
        else:
            # FP32-FMA jet kernel: HVP mode 16 P_w flop per jet row (8 GEMM-equivalents), GRAD mode 6 P_w
            k_flops = (16.0 if dom_mode == MODE_HVP else 6.0) * Pw * R * T_local
            achieved = k_flops / (avg_ms * 1e-3) / 1e12 if d_n else None
            roofline = {
                "kernel": f"pde_kernel<W,PDE,{'MODE_HVP' if dom_mode == MODE_HVP else 'MODE_GRAD'}> (fused SIREN forward jet + adjoint"
                          + (" + tangent-of-adjoint)" if dom_mode == MODE_HVP else ")"),
                "bound": "fp32_fma", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                "frac": (achieved / fp32_peak) if achieved else None,
                "peak_source": "of measured: " + fp32_src,
                "traffic": traffic, "flops_per_launch": k_flops, "avg_launch_ms": avg_ms, "launches_timed": d_n,
                "kernel_share_of_step": share, "per_mode_ms": per_mode,
            }
    flops = step_flops(wl, sizes)
    metric, unit = METRIC[kind]
    grad_ms, grad_n = prof.get(MODE_GRAD, (0.0, 0))
    line = {
        "metric": metric, "value": metric_value(wl, sizes, ms_step), "unit": unit, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if job["shards"] else "replicas", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(wl, name, world),
        "e2e": {"value": metric_value(wl, sizes, ms_e2e), "unit": unit, "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": host[0].nbytes(), "d2h_bytes_per_step": 4 * sum(int(x.numel()) for x in out),
                "api": job["api"] + " (pinned host points in, host losses out)"},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "roofline": roofline,
        "step_tflops": flops / (ms_step * 1e-3) / 1e12,
        "steps_per_sec": 1e3 / ms_step,
        "residual_grad_evals_per_sec": (T_local * sizes[-1] * grad_n) / (grad_ms * 1e-3) * (world if job["shards"] else 1) if grad_n else None,
        "losses_finite": finite,
        "checksum": checksum,
        "l2_flush_between_steps": flush is not None,
    }
    if kind == "finetune" and pde == "poisson" and W == 64 and L == 3 and n == 1024:
        # the one published figure of this path: 0.01599 s per fine-tune step, CPU (src/notebooks/resultAnalysis.ipynb:443)
        line["vs_baseline"] = line["value"] / (sizes[-1] / 0.01599)
        line["published_reference"] = {"s_per_step": 0.01599, "evals_per_sec": sizes[-1] / 0.01599, "hardware": "CPU (not stated)",
                                       "source": "src/notebooks/resultAnalysis.ipynb:443"}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(wl, steps=4, warmup=1, budget_s=40.0)
        line["vs_cpu_baseline"] = line["e2e"]["value"] / line["cpu_baseline"]["value"]
    sys.stdout.flush()
    os.dup2(stdout_fd, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--nosiren", action="store_true",
                    help="the reference's --nosiren network: swish hidden layers, no omega (src/maml_pde.py:115); "
                         "FP32-FMA kernels at every width")
    ap.add_argument("--flush-l2", action="store_true", help="write 256 MB between timed steps (small, L2-resident configs)")
    ap.add_argument("--graph", action="store_true",
                    help="replay the MAML rollout as one CUDA graph (launch-bound small configs; per-kernel timing is then off)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload], act="swish" if args.nosiren else "sin")
    bench_config.graph = args.graph
    if args.impl == "reference":
        run_reference(args, wl, args.workload)
    else:
        run_ours(args, wl, args.workload)


if __name__ == "__main__":
    main()
