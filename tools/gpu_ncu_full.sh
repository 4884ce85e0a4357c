#!/bin/bash
# ncu --set full of one launch of a tensor-core kernel at C5 size; exports the raw and source pages as gzipped CSV
# (the .ncu-rep with imported source is too large for gpurun_out and is deleted).
# usage: gpu_ncu_full.sh TAG MODE KERNEL_REGEX [MODE KERNEL_REGEX ...]
set -u
OUT=gpurun_out; mkdir -p $OUT
TAG=$1; shift
while [ $# -ge 2 ]; do
  MODE=$1; KN=$2; shift 2
  timeout 300 python tools/ncu_one.py $MODE 2 > $OUT/${TAG}_${MODE}_${KN}_plain.log 2>&1 || { echo "plain run failed"; tail -3 $OUT/${TAG}_${MODE}_${KN}_plain.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KN -s 1 -c 1 -f -o /tmp/${TAG}_${MODE}_$KN python tools/ncu_one.py $MODE 2 > $OUT/${TAG}_${MODE}_${KN}_ncu.log 2>&1
  echo "ncu $KN exit $?"
  ncu -i /tmp/${TAG}_${MODE}_$KN.ncu-rep --page raw --csv 2>/dev/null | gzip > $OUT/${TAG}_${MODE}_${KN}_raw.csv.gz
  ncu -i /tmp/${TAG}_${MODE}_$KN.ncu-rep --page source --csv 2>/dev/null | gzip > $OUT/${TAG}_${MODE}_${KN}_source.csv.gz
  ncu -i /tmp/${TAG}_${MODE}_$KN.ncu-rep --page details 2>/dev/null > $OUT/${TAG}_${MODE}_${KN}_details.txt
  TRAFFIC_ARGS="${TRAFFIC_ARGS:-} ${MODE}_$KN $OUT/${TAG}_${MODE}_${KN}_raw.csv.gz"
  python tools/ncu_source_summary.py $OUT/${TAG}_${MODE}_${KN}_source.csv.gz 25 > $OUT/${TAG}_${MODE}_${KN}_source_summary.txt 2>&1
  python tools/ncu_source_bins.py $OUT/${TAG}_${MODE}_${KN}_source.csv.gz 400 >> $OUT/${TAG}_${MODE}_${KN}_source_summary.txt 2>&1
done
python tools/ncu_traffic.py $TRAFFIC_ARGS > $OUT/${TAG}_traffic.json 2>&1
ls -la $OUT | tail
