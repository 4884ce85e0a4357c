#!/bin/bash
# kernel iteration: tensor-core parity subset + C5-size timings (no pytest)
set -u
TAG=${1:-r02i}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python tools/tc_check.py parity > $OUT/${TAG}_tc_parity.log 2>&1; echo "parity exit $?"; tail -8 $OUT/${TAG}_tc_parity.log | cut -c1-330
timeout 600 python tools/tc_check.py time loss grad hvp > $OUT/${TAG}_tc_time.log 2>&1; echo "time exit $?"; tail -8 $OUT/${TAG}_tc_time.log
