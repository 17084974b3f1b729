#!/bin/bash
# GPU session for the encoder work: new-kernel tests first (bounded), whole GPU suite, micro-benchmarks, bench at B=64.
# Usage: bash tools/gpu_enc.sh <tag>
set -u
TAG=${1:-enc}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_stages.py -x -q -k "persistent or enc_attention or tcgen05" > $OUT/pytest_new.log 2>&1; echo "new-kernel tests exit $?"; tail -15 $OUT/pytest_new.log
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -15 $OUT/pytest_gpu.log
timeout 600 python tools/microbench.py > $OUT/microbench.jsonl 2> $OUT/microbench.err; echo "microbench exit $?"; grep -v wstream $OUT/microbench.jsonl; tail -3 $OUT/microbench.err
timeout 900 python bench.py --batch 64 --steps 2 --warmup 3 > $OUT/bench_b64.json 2> $OUT/bench_b64.err; echo "bench exit $?"; tail -c 2500 $OUT/bench_b64.json; tail -5 $OUT/bench_b64.err
