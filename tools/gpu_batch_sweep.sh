for cfg in "448 4" "512 4"; do set -- $cfg
WB200_LANES=$2 timeout 600 python bench.py --batch $1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_b$1_l$2.json 2> gpurun_out/bench_b$1_l$2.err; echo "B=$1 lanes=$2 exit $?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_b$1_l$2.json"))
    print({k:d[k] for k in ("value","ms_per_step","phase_ms_per_step_host_clock")}, d["e2e"]["value"])
except Exception as e: print("no json", e)
PY
tail -2 gpurun_out/bench_b$1_l$2.err
done
