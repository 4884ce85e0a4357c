#!/bin/bash
# tcgen05 bring-up probe (each stage its own process: a trap in one must not hide the others)
set -u
OUT=gpurun_out; mkdir -p $OUT
for w in 1 2 4 8; do
  timeout 180 ./tools/tc_probe $w > $OUT/probe_$w.log 2>&1; echo "probe $w exit $?"
  cat $OUT/probe_$w.log
done
