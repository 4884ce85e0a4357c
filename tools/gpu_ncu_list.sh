#!/bin/bash
# per-kernel device times of one tc_check timing run (cold-cache, serialised: compare shares)
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/tc_check.py time "$@" > $OUT/ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/tc_launches.csv python tools/tc_check.py time "$@" > $OUT/ncu_list.log 2>&1
echo "ncu exit $?"
python tools/summarize_launches.py $OUT/tc_launches.csv $OUT/tc_launches.md "tc_check time $*" 2>&1 | head -30; rm -f $OUT/tc_launches.csv
