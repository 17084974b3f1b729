#!/bin/bash
# strong scaling of ONE long clip (bench.py --workload clip): BASELINE config 4 (medium, 1 h = 120 shards) and config 5 (large-v3, 8 h = 960
# shards) on N GPUs of one box.  usage: gpu_clip.sh <tag> <N> [configs: "4 5"]
set -u
TAG=${1:-clip}; N=${2:-1}; CFGS=${3:-"4 5"}; OUT=gpurun_out/$TAG; mkdir -p $OUT
run() { # name, bench args...
  name=$1; shift
  if [ "$N" = "1" ]; then
    timeout 1500 python bench.py --gpus 1 --workload clip "$@" > $OUT/$name.json 2> $OUT/$name.err
  else
    timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --workload clip "$@" > $OUT/$name.json 2> $OUT/$name.err
  fi
  echo "== $name N=$N exit $? $(python -c "import json;d=json.loads(open('$OUT/$name.json').read().strip().splitlines()[-1]);print({k:d[k] for k in ('value','ms_per_step','shards_total','shards_per_gpu_max','segments_in_clip_order_with_offsets','n_tokens')}, d['clocks']['sm_mhz'])" 2>&1)"
  tail -2 $OUT/$name.err | cut -c1-200
}
for c in $CFGS; do
  case $c in
    4) run scale_clip_cfg4_medium_1h_n$N --model medium --hours 1 --steps ${STEPS:-3} --warmup 1 ;;
    5) run scale_clip_cfg5_large-v3_8h_n$N --model large-v3 --hours 8 --steps ${STEPS5:-1} --warmup 1 ;;
  esac
done
