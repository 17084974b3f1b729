#!/bin/bash
# quick GPU iteration: step-kernel tests, whole GPU suite, micro-benchmarks, short bench.  Usage: bash tools/gpu_iter.sh <tag> [bench args]
set -u
TAG=${1:-it}; shift || true
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_step_kernels.py -x -q > $OUT/pytest_step.log 2>&1; echo "step-kernel tests exit $?"; tail -5 $OUT/pytest_step.log
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -15 $OUT/pytest_gpu.log
timeout 600 python tools/microbench.py > $OUT/microbench.jsonl 2> $OUT/microbench.err; echo "microbench exit $?"; cat $OUT/microbench.jsonl; tail -3 $OUT/microbench.err
timeout 900 python bench.py "$@" > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; tail -c 2500 $OUT/bench.json; tail -5 $OUT/bench.err
