#!/bin/bash
# quick GPU loop: tensor-core parity cases, then C5-size timings with per-kernel breakdown
set -u
OUT=gpurun_out; mkdir -p $OUT
TAG=${1:-q}; shift || true
timeout 600 python tools/tc_check.py parity > $OUT/${TAG}_parity.log 2>&1; echo "parity exit $?"; cut -c1-400 $OUT/${TAG}_parity.log | tail -8
timeout 600 python tools/tc_check.py time "${@:-grad hvp}" > $OUT/${TAG}_time.log 2>&1; echo "time exit $?"; tail -6 $OUT/${TAG}_time.log
