#!/bin/bash
# One `ncu --set full` capture (1 GPU) of the kernels matching REGEX in CMD, after CMD exited 0 without ncu.
#   bash tools/r2_ncu_full.sh TAG REGEX SKIP COUNT -- CMD...
TAG=$1; REGEX=$2; SKIP=$3; COUNT=$4; shift 5
O=gpurun_out
mkdir -p $O
timeout 600 "$@" > $O/${TAG}_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s $SKIP -c $COUNT \
   -o $O/${TAG}_full -f "$@" > $O/${TAG}_ncu_full.log 2>&1
tail -5 $O/${TAG}_plain.log; tail -3 $O/${TAG}_ncu_full.log
