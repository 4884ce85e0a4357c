#!/bin/bash
set -u
TAG=${1:-r02e}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c5.json 2> $OUT/${TAG}_bench_c5.err; echo "bench c5 exit $?"
cat $OUT/${TAG}_bench_c5.json | cut -c1-3000; tail -3 $OUT/${TAG}_bench_c5.err
