#!/bin/bash
# round 2, final measurement pass on one B200: GPU parity tests, smoke, C1-C5 bench lines (+ the reference arm), the
# ncu launch list of the default bench command, ncu --set full of one launch of each tensor-core kernel at C5 size,
# deployment_time lines.  Everything lands in gpurun_out/${TAG}_*; the summaries are copied to profiles/ by hand.
set -u
TAG=${1:-r02_final}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader; nproc; free -g | head -2
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log; tail -3 $OUT/${TAG}_pytest.log
timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/${TAG}_smoke.log
for WL in c1_poisson_finetune_3x64_1k c2_poisson_leap_3x64_T16_1k c3_burgers_maml_3x64_T8_1k c4_hyper_maml_3x64_T8_1k; do
  timeout 600 python bench.py --workload $WL --steps 20 --warmup 5 > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err; echo "bench $WL exit $?"
done
timeout 900 python bench.py > $OUT/${TAG}_bench_c5.json 2> $OUT/${TAG}_bench_c5.err; echo "bench c5 exit $?"; cut -c1-400 $OUT/${TAG}_bench_c5.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > $OUT/${TAG}_bench_c5_reference_arm.json 2> $OUT/${TAG}_bench_c5_reference_arm.err; echo "reference arm exit $?"
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv $BCMD > $OUT/${TAG}_ncu_list.log 2>&1; echo "ncu list exit $?"
python tools/summarize_launches.py $OUT/${TAG}_launches.csv $OUT/${TAG}_ncu_launches_c5.md "bench.py --steps 1 --warmup 3 --no-cpu-baseline (C5, round-2 kernels)" > /dev/null 2>&1; rm -f $OUT/${TAG}_launches.csv
head -14 $OUT/${TAG}_ncu_launches_c5.md
bash tools/gpu_ncu_full.sh ${TAG}_ncu hvp tc_hvp_kernel grad tc_pde_kernel loss tc_pde_kernel grad tc_wgrad_kernel hvp tc_wgrad_kernel > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
cat $OUT/${TAG}_ncu_traffic.json | head -60
timeout 600 python tools/deploy_time.py > $OUT/${TAG}_deploy_time.jsonl 2> $OUT/${TAG}_deploy_time.err; echo "deploy exit $?"
