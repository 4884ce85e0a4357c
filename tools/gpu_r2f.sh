#!/bin/bash
# round 2: GPU parity tests (incl. the deployment rollout), deployment_time lines, compute-sanitizer memcheck/racecheck
# of smoke() and of one tensor-core MAML parity test
set -u
TAG=${1:-r02f}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
timeout 600 python tools/deploy_time.py > $OUT/${TAG}_deploy_time.jsonl 2> $OUT/${TAG}_deploy_time.err; echo "deploy exit $?"; cat $OUT/${TAG}_deploy_time.jsonl | cut -c1-700; tail -3 $OUT/${TAG}_deploy_time.err
if [ "${SANITIZE:-1}" = "1" ]; then
  for TOOL in memcheck racecheck; do
    timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 python __graft_entry__.py smoke > $OUT/${TAG}_sanitizer_${TOOL}_smoke.log 2>&1; echo "$TOOL smoke exit $?"
    tail -4 $OUT/${TAG}_sanitizer_${TOOL}_smoke.log
    timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "test_maml_outer_c5_network_k5_clip_active and poisson" > $OUT/${TAG}_sanitizer_${TOOL}_maml_tc.log 2>&1; echo "$TOOL maml exit $?"
    tail -4 $OUT/${TAG}_sanitizer_${TOOL}_maml_tc.log
  done
fi
