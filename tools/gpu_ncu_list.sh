#!/bin/bash
# ncu launch list of a bench command (after the same command exited 0 without ncu).  Usage: bash tools/gpu_ncu_list.sh <tag> <skip> <count> [bench args]
set -u
TAG=$1; SKIP=$2; CNT=$3; shift 3
OUT=gpurun_out/$TAG; mkdir -p $OUT
BCMD="python bench.py $*"
$BCMD > $OUT/bench_plain.json 2> $OUT/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c $CNT --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"; tail -c 600 $OUT/bench_plain.json; python tools/ncu_launches.py $OUT/launches.csv | head -30
