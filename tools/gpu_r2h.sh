#!/bin/bash
# round 2, pass h: branch-free sincos (compat.cuh) -- GPU parity suite, then the same timings as pass g / f
mkdir -p gpurun_out
(timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/r02_h_gpu_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_h_gpu_pytest.log)
tail -n 3 gpurun_out/r02_h_gpu_pytest.log
python - <<'P' > gpurun_out/r02_h_kernel_time_3x64.jsonl 2>&1
import json, os, sys, torch
sys.path.insert(0, os.getcwd())
from meta_pde_b200 import ops, synthetic
K = ops.kernels(torch.device("cuda:0")); dev = K.device
for (W, L, T, n) in ((64, 3, 64, 16384), (64, 3, 1, 1024), (256, 5, 8, 16384)):
    sets, prm = synthetic.make_tasks("poisson", T, [n, n], seed=1); sets = [s.to(dev) for s in sets]; prm = prm.to(dev)
    th = synthetic.siren_theta([W] * L, 1, 30.0, 30.0, True, seed=0).to(dev).unsqueeze(0).repeat(T, 1).contiguous(); v = torch.randn_like(th)
    net = K.net_desc(L, W, 1, 30.0, True)
    fns = {"loss": lambda: K.loss(net, "poisson", 1.0, th, sets, prm, T), "grad": lambda: K.loss_grad(net, "poisson", 1.0, th, sets, prm, T),
           "hvp": lambda: K.loss_hvp(net, "poisson", 1.0, th, v, sets, prm, T)}
    for mode, fn in fns.items():
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): fn()
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"act": "sin", "mode": mode, "W": W, "L": L, "T": T, "n": n, "ms_per_call": e0.elapsed_time(e1) / 10}))
P
cat gpurun_out/r02_h_kernel_time_3x64.jsonl
for w in c1_poisson_finetune_3x64_1k c3_burgers_maml_3x64_T8_1k c2_poisson_leap_3x64_T16_1k c4_hyper_maml_3x64_T8_1k; do
  timeout 60 python bench.py --workload $w --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r02_h_bench_$w.json 2> gpurun_out/r02_h_bench_$w.err; echo "bench $w exit $?"
done
timeout 150 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_h_bench_c5.json 2> gpurun_out/r02_h_bench_c5.err; echo "bench c5 exit $?"
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_h_bench_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); r = d.get("roofline") or {}
        print(f.split("bench_")[1], round(d["value"], 4), "ms/step", round(d["ms_per_step"], 4), "frac", r.get("frac"), r.get("frac_issued"), r.get("per_mode_ms"))
    except Exception as e:
        print(f, "unreadable", e)
P
