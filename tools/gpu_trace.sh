#!/bin/bash
# in-situ step timelines at B=256 for a few lane configurations (one bench step each)
set -u
OUT=gpurun_out/${1:-trace}; mkdir -p $OUT
run() { # tag, env...
  tag=$1; shift
  rm -f $OUT/trace_$tag.jsonl
  env "$@" WB200_STEP_TRACE=$OUT/trace_$tag.jsonl timeout 600 python bench.py --batch ${BATCH:-256} --steps 1 --warmup 1 --no-cpu-baseline > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
  echo "== $tag exit $? $(python -c "import json;d=json.load(open('$OUT/bench_$tag.json'));print(d['ms_per_step'], d['phase_ms_per_step_host_clock']['decode_step'])" 2>&1)"
  python tools/trace_summary.py $OUT/trace_$tag.jsonl | tail -${TAILN:-4}
}
