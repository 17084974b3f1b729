#!/bin/bash
# in-situ step timelines at B=256 for a few lane configurations (one bench step each)
set -u
OUT=gpurun_out/${1:-trace}; mkdir -p $OUT
run() { # tag, env...
  tag=$1; shift
  rm -f $OUT/trace_$tag.jsonl
  env "$@" WB200_STEP_TRACE=$OUT/trace_$tag.jsonl timeout 600 python bench.py --batch ${BATCH:-256} --steps 1 --warmup 1 --no-cpu-baseline --no-extras --api ${API:-batch} > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
  echo "== $tag exit $? $(python -c "import json;d=json.load(open('$OUT/bench_$tag.json'));print(d['ms_per_step'], d['phase_ms_per_step_host_clock']['decode_step'])" 2>&1)"
  python tools/trace_summary.py $OUT/trace_$tag.jsonl | tail -${TAILN:-4}
}
for cfg in ${CFGS:-"lanes4"}; do
  case $cfg in
    lanes4) run lanes4 WB200_LANES=4 ;;
    lanes2) run lanes2 WB200_LANES=2 ;;
    lanes1) run lanes1 WB200_LANES=1 ;;
    lanes4_nopdl) run lanes4_nopdl WB200_LANES=4 WB200_NO_PDL=1 ;;
    lanes3) run lanes3 WB200_LANES=3 ;;
    lanes5) run lanes5 WB200_LANES=5 ;;
    lanes6) run lanes6 WB200_LANES=6 ;;
    cps1) run cps1 WB200_XS_CPS=1 WB200_XS_STAGES=6 ;;
    wc40) run wc40 WB200_WC_CTAS=40 ;;
    wc74) run wc74 WB200_WC_CTAS=74 ;;
    noprio) run noprio WB200_NO_PRIO=1 ;;
  esac
done
