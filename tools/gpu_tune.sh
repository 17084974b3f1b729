#!/bin/bash
# env-knob sweep of the decode lanes at the bench batch: one short bench per configuration ("name:VAR=val,VAR=val")
set -u
OUT=gpurun_out/${1:-tune}; mkdir -p $OUT; shift
for cfg in "$@"; do
  name=${cfg%%:*}; envs=${cfg#*:}; envs=${envs//,/ }
  env $envs timeout 600 python bench.py --steps ${STEPS:-2} --warmup 1 --no-cpu-baseline --no-extras --api ${API:-batch} --batch ${BATCH:-384} > $OUT/bench_$name.json 2> $OUT/bench_$name.err
  echo "== $name [$envs] exit $? $(python -c "import json;d=json.load(open('$OUT/bench_$name.json'));print(round(d['value'],1), round(d['ms_per_step'],1), d['phase_ms_per_step_host_clock']['encode'], d['decode_ms_per_30s_window'], d['clocks']['sm_mhz'])" 2>&1)"
done
