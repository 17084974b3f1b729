#!/bin/bash
# GPU session for the decode-step work: step-kernel tests (bounded), whole GPU suite, micro-benchmarks, bench at B=64 and B=256
set -u
TAG=${1:-step}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_step_kernels.py -x -q > $OUT/pytest_step.log 2>&1; echo "step-kernel tests exit $?"; tail -8 $OUT/pytest_step.log
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -12 $OUT/pytest_gpu.log
timeout 600 python tools/microbench.py > $OUT/microbench.jsonl 2> $OUT/microbench.err; echo "microbench exit $?"; grep "wstream\|wcluster" $OUT/microbench.jsonl; tail -3 $OUT/microbench.err
for b in 64 256; do
  timeout 900 python bench.py --batch $b --steps 2 --warmup 2 --no-cpu-baseline > $OUT/bench_b$b.json 2> $OUT/bench_b$b.err; echo "bench B=$b exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_b$b.json"))
    print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock","gpu_launches")})
except Exception as e:
    print("no bench json:", e)
PY
  tail -3 $OUT/bench_b$b.err
done
