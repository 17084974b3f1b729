timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
WB200_HOST_PROF=1 timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_v7.json 2> gpurun_out/bench_v7.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_v7.json"))
print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock")}, d["e2e"]["value"])
PY
grep "host_prof: total" gpurun_out/bench_v7.err | tail -2 | cut -c1-600
