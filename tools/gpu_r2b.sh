#!/bin/bash
# round 2, kernel iteration: tensor-core parity subset + C5-size timings of the three modes
set -u
TAG=${1:-r02b}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python tools/tc_check.py parity > $OUT/${TAG}_tc_parity.log 2>&1; echo "parity exit $?"; tail -8 $OUT/${TAG}_tc_parity.log | cut -c1-400
timeout 600 python tools/tc_check.py time loss grad hvp > $OUT/${TAG}_tc_time.log 2>&1; echo "time exit $?"; tail -8 $OUT/${TAG}_tc_time.log
timeout 1500 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
