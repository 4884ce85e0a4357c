#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the C5 bench line (tensor-core path), ncu launch list of the same
# command, ncu --set full of one launch each of tc_hvp_kernel / tc_pde_kernel(GRAD) / tc_wgrad_kernel at C5.
set -u
TAG=${1:-r01d}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/${TAG}_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv $BCMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?"
python tools/summarize_launches.py $OUT/${TAG}_launches.csv $OUT/${TAG}_launches.md "bench.py --steps 1 --warmup 3 (C5, tensor-core path)" > /dev/null 2>&1
head -25 $OUT/${TAG}_launches.md; rm -f $OUT/${TAG}_launches.csv
for KN in tc_hvp_kernel tc_pde_kernel tc_wgrad_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KN -s 12 -c 1 -o $OUT/${TAG}_full_$KN $BCMD > $OUT/${TAG}_ncu_full_$KN.log 2>&1
  echo "ncu full $KN exit $?"
done
ls -la $OUT | tail -14
