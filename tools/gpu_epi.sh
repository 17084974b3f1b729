#!/bin/bash
set -u
TAG=${1:-epi}; OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in 1 4 5 6; do
  WB200_EPI=$v timeout 300 python tools/microbench_gemm.py 64 >> $OUT/gemm_variants.jsonl 2>> $OUT/err.log; echo "variant $v exit $?"
done
grep "epi_only\|o_proj" $OUT/gemm_variants.jsonl | cut -c1-220
