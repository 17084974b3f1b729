#!/bin/bash
# round-2 sanity: smoke, whole GPU suite, one short bench through the scheduler
set -u
TAG=${1:-r2_check}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -15 $OUT/pytest_gpu.log
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("$OUT/bench.json"))
print({k:d.get(k) for k in ("value","ms_per_step","phase_ms_per_step_host_clock")}, d["e2e"]["value"])
r=d["roofline"]; print({k:r.get(k) for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est","traffic")})
for r in d.get("roofline_next",[]): print({k:r.get(k) for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est")})
PY
tail -5 $OUT/bench.err
