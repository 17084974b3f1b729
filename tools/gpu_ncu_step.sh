#!/bin/bash
# ncu launch list of the decode-step kernels at a given batch (graph nodes are profiled individually; cold-cache, serialised)
set -u
TAG=$1; B=${2:-64}; OUT=gpurun_out/$TAG; mkdir -p $OUT
BCMD="python bench.py --batch $B --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $BCMD > $OUT/bench_plain.json 2> $OUT/bench_plain.err || { echo "plain bench failed"; tail -5 $OUT/bench_plain.err; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ln_reduce|cross_attn_step|gemm_wstream|self_attn_step|gelu_reduce|sampler" -s 2000 -c 1200 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"; python tools/ncu_launches.py $OUT/launches.csv | head -30
