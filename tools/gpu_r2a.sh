#!/bin/bash
# round 2, call A: GPU parity tests (incl. the new train-step / tensor-core rollout / 16k-point oracle tests), smoke,
# bench lines of C1-C4 and the default C5 line.
set -u
TAG=${1:-r02a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
free -g | head -2; nproc
timeout 2400 python -m pytest tests -m gpu -q -x --durations=15 > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -30 $OUT/${TAG}_pytest.log
timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/${TAG}_smoke.log
for WL in c1_poisson_finetune_3x64_1k c2_poisson_leap_3x64_T16_1k c3_burgers_maml_3x64_T8_1k c4_hyper_maml_3x64_T8_1k; do
  timeout 600 python bench.py --workload $WL --steps 20 --warmup 5 > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err; echo "bench $WL exit $?"
  cut -c1-600 $OUT/${TAG}_bench_$WL.json; tail -2 $OUT/${TAG}_bench_$WL.err
done
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_c5.json 2> $OUT/${TAG}_bench_c5.err; echo "bench c5 exit $?"
cat $OUT/${TAG}_bench_c5.json; tail -3 $OUT/${TAG}_bench_c5.err
