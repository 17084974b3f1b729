#!/bin/bash
# ncu --set full of every kernel the hot path names, one capture per kernel from the SHIPPING build at the bench's per-lane shapes
# (large-v3, one lane of 112 sequences = a lane of the default 448-shard bench step), plus the DTW kernels from the alignment test.  Summaries -> $OUT/<name>.txt
set -u
OUT=gpurun_out/${1:-ncu_all}; mkdir -p $OUT
BCMD="python bench.py --batch 112 --steps 1 --warmup 1 --no-extras --no-cpu-baseline --api batch"
DCMD="python -m pytest tests/test_dtw.py -q -x -k alignment_chain"
cap() { # name regex skip count cmd...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -f -o $OUT/$name "$@" > $OUT/$name.log 2>&1
  rc=$?
  if [ -f $OUT/$name.ncu-rep ]; then
    ncu -i $OUT/$name.ncu-rep --page raw --csv > $OUT/$name.raw.csv 2>/dev/null
    python tools/ncu_raw.py $OUT/$name.raw.csv > $OUT/$name.txt 2>&1
    echo "== $name rc=$rc $(grep -c -- '----' $OUT/$name.txt) launch(es)"; grep -E "Kernel Name|time_duration|dram__bytes_read|tensor_cycles|dram_throughput" $OUT/$name.txt | head -12
    rm -f $OUT/$name.raw.csv
  else
    echo "== $name rc=$rc NO REPORT"; tail -3 $OUT/$name.log
  fi
}
cap gemm_f16_persist  "gemm_f16_persist"        700 4 $BCMD
cap enc_attn          "enc_attn_kernel"          40 1 $BCMD
cap layernorm_rows    "layernorm_kernel"        100 1 $BCMD
cap gelu_table        "gelu_table_kernel"        40 1 $BCMD
cap mel_frames        "mel_frames_kernel"       100 1 $BCMD
cap mel_window        "mel_window_kernel|mel_normalize_kernel" 100 2 $BCMD
cap gemm_wcluster     "gemm_wcluster_kernel"   3000 6 $BCMD
cap self_attn_step    "self_attn_step_kernel"  1000 1 $BCMD
cap ln_reduce         "ln_reduce_kernel"       2000 1 $BCMD
cap cross_attn_step   "cross_attn_step_kernel" 1000 1 $BCMD
cap logits_wstream    "gemm_wstream_kernel"     100 1 $BCMD
cap sampler           "sampler_kernel"          100 1 $BCMD
cap dtw               "dtw_wavefront_kernel|dtw_norm_kernel|dtw_medfilt_mean_kernel" 0 3 $DCMD
ls -la $OUT/*.ncu-rep | awk '{print $5, $9}'
