#!/bin/bash
# round 2, N-GPU box: 1-GPU == 2-GPU hardware test, then bench.py at N = 1 .. NG (torchrun), default workload (C5)
set -u
NG=${1:-2}; TAG=${2:-r02_multi}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $OUT/${TAG}_pytest_multi.log 2>&1; echo "multi pytest exit $?"; tail -3 $OUT/${TAG}_pytest_multi.log
for N in 1 2 4 8; do
  [ $N -gt $NG ] && break
  if [ $N -eq 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c5_n1.json 2> $OUT/${TAG}_bench_c5_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c5_n$N.json 2> $OUT/${TAG}_bench_c5_n$N.err
  fi
  echo "bench N=$N exit $?"; python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/${TAG}_bench_c5_n$N.json") if l.startswith("{")][-1])
    print("N=$N", d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d.get("checksum"))
except Exception as e:
    print("parse failed", e)
PY
done
