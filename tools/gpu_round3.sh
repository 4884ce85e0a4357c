#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the C5 bench line, ncu --set full of the HVP jet kernel (dev size + C5).
set -u
TAG=${1:-r01c}
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cat $OUT/${TAG}_bench.json
for WL in dev_poisson_maml_5x256_T8_2k c5_poisson_maml_5x256_T64_16k; do
  timeout 600 python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain_$WL.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
      -k 'regex:pde_kernel<.*\(int\)2>' -s 5 -c 1 -o $OUT/${TAG}_hvp_$WL python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline \
      > $OUT/${TAG}_ncu_full_$WL.log 2>&1
  echo "ncu full $WL exit $?"
done
ls -la $OUT | tail -12
