#!/bin/bash
# GPU parity tests + ncu full capture of the HVP jet kernel (dev size first, then the C5 bench command).
set -u
TAG=${1:-r01b}
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
for WL in dev_poisson_maml_5x256_T8_2k c5_poisson_maml_5x256_T64_16k; do
  timeout 600 python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain_$WL.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
      -k 'regex:pde_kernel<.*\(int\)2>' -s 5 -c 1 -o $OUT/${TAG}_hvp_$WL python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline \
      > $OUT/${TAG}_ncu_full_$WL.log 2>&1
  echo "ncu full $WL exit $?"
done
ls -la $OUT | tail -12
