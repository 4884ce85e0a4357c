#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python tools/tc_check.py parity > $OUT/tc_parity.log 2>&1; echo "parity exit $?"; tail -12 $OUT/tc_parity.log | cut -c1-1200
timeout 300 python tools/tc_check.py time loss grad > $OUT/tc_time.log 2>&1; echo "time exit $?"; tail -6 $OUT/tc_time.log
MPDE_DISABLE_TC=1 timeout 300 python tools/tc_check.py time grad > $OUT/ffma_time.log 2>&1; echo "time exit $?"; tail -4 $OUT/ffma_time.log
