#!/bin/bash
# quick check of a decode-step change: step-kernel tests, whole GPU suite, cluster GEMM micro-benchmarks, one short bench
set -u
TAG=${1:-quick}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_step_kernels.py -x -q > $OUT/pytest_step.log 2>&1; echo "step-kernel tests exit $?"; tail -4 $OUT/pytest_step.log
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -6 $OUT/pytest_gpu.log
timeout 600 python tools/microbench.py > $OUT/microbench.jsonl 2> $OUT/microbench.err; echo "microbench exit $?"; grep "wcluster" $OUT/microbench.jsonl
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("$OUT/bench.json"))
print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock")}, d["e2e"]["value"])
r=d["roofline"]; print({k:r[k] for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est","traffic")})
for r in d["roofline_next"]: print({k:r[k] for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est")})
PY
tail -3 $OUT/bench.err
