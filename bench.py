#!/usr/bin/env python
"""bench.py -- MAML outer steps/s on BASELINE.json's scaled Poisson config (C5), 1..8 B200s.

    python bench.py --gpus N --steps K --warmup W          (N > 1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                   (the CPU arm: the oracle port on the host cores)

One "step" = one full outer step of src/maml_pde.py:203-226 over the whole meta-batch: for each of the
64 tasks K=5 inner SGD steps (loss+grad), K outer loss+grads, K Hessian-vector products, 2 loss-only evals
(SURVEY.md 9.4), the task mean (NCCL all-reduce of [g_theta | g_lrs] when N > 1), the clip and the two Adam
updates.  Nothing is skipped or cached inside the timed region.  The 64 tasks are sharded over the ranks
(64/N tasks per GPU: "strong" scaling, the config fixes the global meta-batch).

Printed JSON line (rank 0): see the task contract; `value` = steps/s with this step's points already in HBM,
`e2e` = the same step through MamlTrainer.train_step_host with pinned HOST buffers (H2D of the step's points
and D2H of its losses inside the timed region), `roofline` = the dominant kernel (the fused jet kernel in HVP
mode) timed live with CUDA events on its launch stream, `cpu_baseline` = the oracle timed on the host cores
on a bounded sample (rank 0, N = 1 only).
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
    "c5_poisson_maml_5x256_T64_16k": dict(pde="poisson", L=5, W=256, T=64, n=16384, K=5),
    # small shapes for smoke runs of this script (never the reported configuration)
    "dev_poisson_maml_3x64_T16_1k": dict(pde="poisson", L=3, W=64, T=16, n=1024, K=5),
    "dev_poisson_maml_5x256_T8_2k": dict(pde="poisson", L=5, W=256, T=8, n=2048, K=5),
}
DEFAULT_WORKLOAD = "c5_poisson_maml_5x256_T64_16k"
HYPER = dict(inner_lr=1e-4, outer_lr=1e-3, lr_inner_lr=0.5, inner_grad_clip=100.0, grad_clip=100.0,
             outer_loss_decay=0.1, bc_weight=1.0, omega=30.0)   # pipeline/pipleline_maml_meta.sh:4 values


def weight_count(W, L, out=1):
    return 2 * W + (L - 1) * W * W + W * out


def jet_rows(pde, sizes):
    C = 3 if pde == "hyper_elasticity" else 4
    return C * sizes[-1] + sum(sizes[:-1])


def step_flops(wl, sizes):
    """Algorithmic flops of one outer step over all T tasks: (15K+2) * 2 * P_w * R per task."""
    return (15 * wl["K"] + 2) * 2.0 * weight_count(wl["W"], wl["L"]) * jet_rows(wl["pde"], sizes) * wl["T"]


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
def cpu_baseline(wl, steps=1, warmup=0, sample_tasks=1, sample_n=None, budget_s=25.0):
    """The oracle (oracle/: torch.func restatement of the reference's JAX math, float32 like the reference) timed
    on the host cores on a bounded sample of the workload: `sample_tasks` tasks x `sample_n` boundary+domain
    points of the SAME network / K / second-order outer gradient.  steps/s is scaled to the full meta-batch by
    (tasks * points) proportionality, which is exact for this path (cost is linear in both)."""
    import torch
    from meta_pde_b200 import synthetic
    from oracle import meta as ometa
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_params, oracle_points

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pde, L, W, K = wl["pde"], wl["L"], wl["W"], wl["K"]
    n = sample_n or min(wl["n"], 1024)
    lf = ometa.make_loss_fn(pde, ometa.NetCfg([W] * L, HYPER["omega"], True, 1), HYPER["bc_weight"])
    cfg = ometa.MamlCfg(K, HYPER["inner_lr"], HYPER["inner_grad_clip"], True, HYPER["outer_loss_decay"])
    theta = synthetic.siren_theta([W] * L, 1, HYPER["omega"], HYPER["omega"], True, seed=0)
    lrs = torch.ones(K, theta.numel())
    sizes = synthetic.default_set_sizes(pde, n)
    f32 = torch.float32

    def one_step(seed):
        tasks = []
        prm = synthetic.make_tasks(pde, sample_tasks, sizes, seed=seed)[1]
        mk = lambda s: synthetic.make_tasks(pde, sample_tasks, sizes, seed=s)[0]
        inner = [mk(seed + 10 * k + 1) for k in range(K)]
        outer = [mk(seed + 10 * k + 2) for k in range(K)]
        final, aux = mk(seed + 1001), mk(seed + 1002)
        for t in range(sample_tasks):
            tasks.append(ometa.MamlTask(oracle_params(pde, prm, t, f32), [oracle_points(pde, s, t, f32) for s in inner],
                                        [oracle_points(pde, s, t, f32) for s in outer], oracle_points(pde, final, t, f32),
                                        oracle_points(pde, aux, t, f32)))
        t0 = time.perf_counter()
        ometa.maml_multi_task_grad_and_losses(cfg, lf, theta, tasks, lrs)
        return time.perf_counter() - t0

    for i in range(warmup):
        one_step(100 + i)
    times = []
    for i in range(steps):
        times.append(one_step(200 + i))
        if sum(times) > budget_s:
            break
    t_sample = sum(times) / len(times)
    scale = (wl["T"] / sample_tasks) * (wl["n"] / n)
    return {"value": 1.0 / (t_sample * scale), "unit": "steps/s", "cores": cores, "kind": "port",
            "sample": f"{sample_tasks} task(s) x {n}+{n} points of the {L}x{W} net, K={K}, full second-order outer "
                      f"gradient, float32 torch.func oracle; {len(times)} repetition(s), {t_sample:.2f} s each; scaled "
                      f"x{scale:g} (tasks x points, linear) to the {wl['T']}-task x {wl['n']}-point step",
            "sample_seconds": t_sample, "steps_timed": len(times)}


def run_reference(args, wl, name):
    """--impl reference: the reference's CPU implementation of the path.  JAX/Flax are not installable here
    (DESIGN.md), so this is the oracle port, all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_baseline(wl, steps=max(1, args.steps), warmup=min(args.warmup, 1), sample_n=min(wl["n"], 4096), budget_s=120.0)
    line = {"impl": "reference", "metric": "maml_outer_steps_per_sec", "value": cb["value"], "unit": "steps/s",
            "n_gpus": args.gpus, "steps": cb["steps_timed"], "warmup": min(args.warmup, 1),
            "ms_per_step": 1e3 / cb["value"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(wl, name, args.gpus),
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def bench_config(wl, name, n_gpus):
    return {"workload": name, "pde": wl["pde"], "siren": f"{wl['L']}x{wl['W']} omega=30", "meta_batch_tasks": wl["T"],
            "tasks_per_gpu": wl["T"] // n_gpus, "points_per_set": wl["n"], "inner_steps": wl["K"],
            "second_order": True, "parallelism": f"task-sharded dp{n_gpus}", "cuda_graph": bool(getattr(bench_config, "graph", False)),
            "l2_policy": "inputs larger than L2 (fresh point sets every step; per-step working set >> 126 MB)"}


# ------------------------------------------------------------------------------------------------ GPU arm
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
    from meta_pde_b200 import synthetic, train
    from meta_pde_b200._abi import MODE_GRAD, MODE_HVP, MODE_LOSS

    pde, L, W, T, n, K = (wl[k] for k in ("pde", "L", "W", "T", "n", "K"))
    if T % world:
        raise SystemExit(f"meta-batch {T} does not divide over {world} ranks")
    T_local = T // world
    theta0 = synthetic.siren_theta([W] * L, 1, HYPER["omega"], HYPER["omega"], True, seed=0)
    tr = train.MamlTrainer(pde, L, W, theta0, bsize=T, inner_steps=K, inner_lr=HYPER["inner_lr"], outer_lr=HYPER["outer_lr"],
                           lr_inner_lr=HYPER["lr_inner_lr"], siren_omega=HYPER["omega"], bc_weight=HYPER["bc_weight"],
                           inner_grad_clip=HYPER["inner_grad_clip"], grad_clip=HYPER["grad_clip"],
                           outer_loss_decay=HYPER["outer_loss_decay"], device=dev, use_cuda_graph=args.graph)
    Kn = tr.K
    sizes = synthetic.default_set_sizes(pde, n)

    # synthetic pinned host batches: two alternating point draws (fresh points each step, as the reference samples)
    host = [train.synthetic_maml_batch(pde, T_local, K, n, n, seed=s, pin=True, task_offset=rank * T_local) for s in (1, 2)]
    devb = [hb.to(dev) for hb in host]
    torch.cuda.synchronize()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        sync_all()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    step_no = [0]

    def dev_step(i):
        tr.train_step(step_no[0], devb[i % 2])
        step_no[0] += 1

    last = {}

    def host_step(i):
        last["out"] = tr.train_step_host(step_no[0], host[i % 2])
        step_no[0] += 1

    # ---- warm-up, then EXACTLY `steps` timed steps with the points resident in HBM --------------------------
    for i in range(args.warmup):
        dev_step(i)
    Kn.prof_enable(not args.graph)          # event pairs cannot be recorded into a replayed graph
    launches0 = Kn.launch_count()
    with ClockSampler(local) as clk:
        ms_total = timed(dev_step, args.steps)
    launches = Kn.launch_count() - launches0 + (tr.graph_launches * args.steps if args.graph else 0)
    prof = {m: Kn.prof_read(m) for m in (MODE_LOSS, MODE_GRAD, MODE_HVP)}
    wgrad_ms, wgrad_n = Kn.prof_read(3)
    Kn.prof_enable(False)
    ms_step = ms_total / args.steps
    # ---- end to end: pinned host points -> H2D -> step -> D2H of the losses -------------------------------
    host_step(0)
    ms_e2e = timed(host_step, args.steps) / args.steps
    losses = last["out"][0]
    finite = bool(torch.isfinite(losses).all())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (the fused jet kernel in HVP mode, no gradient accumulation) --------
    Pw, R = weight_count(W, L), jet_rows(pde, sizes)
    hvp_ms, hvp_n = prof[MODE_HVP]
    grad_ms, grad_n = prof[MODE_GRAD]
    loss_ms, loss_n = prof[MODE_LOSS]
    hvp_flops = 16.0 * Pw * R * T_local            # 8 GEMM-equivalents: fwd 1 + tangent 2 + bwd 1 + tangent-bwd 4
    achieved = hvp_flops / (hvp_ms / max(hvp_n, 1) * 1e-3) / 1e12 if hvp_n else None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    fp32_peak, fp32_src = 71.6, "profiles/r01_microbench_pipes.jsonl (FFMA microbench on this pool's B200, 296 CTAs)"
    tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    tc = W in (128, 256) and os.environ.get("MPDE_DISABLE_TC", "0") != "1"
    # dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed `ncu --set full` capture of the same
    # kernel at this workload (profiles/ncu_traffic.json, written by tools/gpu_ncu_full.sh); null for other workloads
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if tj.get("workload") == name and world == 1:
            traffic = tj["kernels"].get("tc_hvp_kernel" if tc else "pde_kernel_hvp", {}).get("traffic_bytes")
    except (OSError, ValueError, KeyError):
        pass
    share = {"hvp": hvp_ms / ms_total, "grad": grad_ms / ms_total, "loss": loss_ms / ms_total, "tc_wgrad": wgrad_ms / ms_total}
    per_mode = {"hvp": [hvp_ms, hvp_n], "grad": [grad_ms, grad_n], "loss": [loss_ms, loss_n], "tc_wgrad": [wgrad_ms, wgrad_n]}
    if tc:
        # tc_hvp_kernel runs 6 of the 8 GEMM-equivalents of a Hessian-vector product (forward 1, tangent forward 2,
        # adjoint 1, tangent adjoint 2) as fp16 hi/lo 3-term splits on tcgen05: 3 tensor-core flops per algorithmic flop.
        # The 2 weight-gradient GEMM-equivalents are tc_wgrad_kernel's.
        hvp_flops = 12.0 * Pw * R * T_local
        achieved = hvp_flops / (hvp_ms / max(hvp_n, 1) * 1e-3) / 1e12 if hvp_n else None
        roofline = {
            "kernel": "tc_hvp_kernel<W,PDE> (tcgen05 kind::f16, fp16 hi/lo 3-term split, TMEM accumulators)",
            "bound": "tensor", "achieved": achieved, "peak": tensor_peak, "unit": "TFLOP/s",
            "frac": (achieved / tensor_peak) if achieved else None,
            "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)",
            "achieved_issued_tflops": (3.0 * achieved) if achieved else None,
            "frac_issued": (3.0 * achieved / tensor_peak) if achieved else None,
            "note": "achieved = ALGORITHMIC fp32 flops; each is 3 fp16 tensor-core flops (a_hi*b_hi + a_lo*b_hi + a_hi*b_lo), "
                    "so frac_issued = 3*frac is the tensor-pipe occupancy and 1/3 is the ceiling of frac at fp32 accuracy",
            "traffic": traffic, "flops_per_launch": hvp_flops, "avg_launch_ms": hvp_ms / max(hvp_n, 1), "launches_timed": hvp_n,
            "kernel_share_of_step": share, "per_mode_ms": per_mode,
        }
    else:
        roofline = {
            "kernel": "pde_kernel<W,PDE,MODE_HVP> (fused SIREN forward jet + adjoint + tangent-of-adjoint)",
            "bound": "fp32_fma", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": (achieved / fp32_peak) if achieved else None,
            "peak_source": "of measured: " + fp32_src,
            "tensor_peak_bf16_sustained": tensor_peak,
            "frac_of_tensor_peak": (achieved / tensor_peak) if achieved else None,
            "traffic": traffic,
            "flops_per_launch": hvp_flops, "avg_launch_ms": hvp_ms / max(hvp_n, 1), "launches_timed": hvp_n,
            "kernel_share_of_step": share, "per_mode_ms": per_mode,
        }
    flops = step_flops(wl, sizes)
    line = {
        "metric": "maml_outer_steps_per_sec", "value": 1e3 / ms_step, "unit": "steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(wl, name, world),
        "e2e": {"value": 1e3 / ms_e2e, "unit": "steps/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": host[0].nbytes(), "d2h_bytes_per_step": 4 * sum(int(x.numel()) for x in last["out"]),
                "api": "meta_pde_b200.train.MamlTrainer.train_step_host (pinned host points in, host losses out)"},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "roofline": roofline,
        "step_tflops": flops / (ms_step * 1e-3) / 1e12,
        "residual_grad_evals_per_sec": (T_local * sizes[-1] * grad_n) / (grad_ms * 1e-3) * world if grad_n else None,
        "losses_finite": finite,
    }
    if world == 1 and not args.no_cpu_baseline:
        # bounded sample: 1 task x 4096+4096 points, 4 timed repetitions after a warm-up (10-20 s on the box's cores)
        line["cpu_baseline"] = cpu_baseline(wl, steps=4, warmup=1, sample_n=min(wl["n"], 4096), budget_s=40.0)
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
    ap.add_argument("--graph", action="store_true",
                    help="replay the rollout as one CUDA graph (launch-bound small configs; per-kernel timing is then off)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    bench_config.graph = args.graph
    if args.impl == "reference":
        run_reference(args, wl, args.workload)
    else:
        run_ours(args, wl, args.workload)


if __name__ == "__main__":
    main()
