#!/bin/bash
# decode-lane sweep: step-kernel + lane parity tests, cross-attention sweep, then bench at a given batch with WB200_LANES = 1..4
set -u
TAG=${1:-lanes}; B=${2:-256}; LANES=${3:-"1 2 3 4"}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_step_kernels.py -x -q > $OUT/pytest_step.log 2>&1; echo "step-kernel tests exit $?"; tail -5 $OUT/pytest_step.log
timeout 600 python -m pytest tests/test_gpu_golden.py -x -q -k "lanes or batch" > $OUT/pytest_lanes.log 2>&1; echo "lane tests exit $?"; tail -8 $OUT/pytest_lanes.log
timeout 600 python tools/microbench_cross.py > $OUT/microbench_cross.jsonl 2> $OUT/microbench_cross.err; echo "microbench exit $?"; cat $OUT/microbench_cross.jsonl; tail -3 $OUT/microbench_cross.err
for k in $LANES; do
  WB200_LANES=$k timeout 900 python bench.py --batch $B --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench_l$k.json 2> $OUT/bench_l$k.err; echo "bench lanes=$k exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_l$k.json"))
    print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock","gpu_launches")}, d["e2e"]["value"])
except Exception as e:
    print("no bench json:", e)
PY
  tail -3 $OUT/bench_l$k.err
done
