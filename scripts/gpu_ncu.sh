#!/bin/bash
# ncu --set full capture of selected kernels.  Usage: gpurun -- 'bash scripts/gpu_ncu.sh tag "regex" skip count'
TAG=${1:-n}; RE=${2:-nms_image}; SKIP=${3:-6}; CNT=${4:-2}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --quick --no-e2e --sets 4"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_$TAG.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/ncu_$TAG.log
