for cfg in "384 4" "384 3" "320 4"; do set -- $cfg
WB200_LANES=$2 timeout 900 python bench.py --batch $1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_b$1_l$2.json 2> gpurun_out/bench_b$1_l$2.err; echo "B=$1 lanes=$2 exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_b$1_l$2.json"))
print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock")}, d["e2e"]["value"])
r=d["roofline"]; print({k:r[k] for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est","event_bracket_overhead_ms")})
for r in d["roofline_next"]: print({k:r[k] for k in ("kernel","achieved","frac","avg_launch_ms","share_of_step_est")})
PY
tail -2 gpurun_out/bench_b$1_l$2.err
done
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
