#!/bin/bash
# ncu --set full capture of the fused encoder attention (micro-benchmark shapes), raw metrics + hottest source lines
set -u
OUT=gpurun_out/${1:-ncu_attn}; mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:enc_attn -s 2 -c 1 -o $OUT/enc_attn_full python tools/microbench.py > $OUT/ncu.log 2>&1
echo "ncu exit $?"
ncu -i $OUT/enc_attn_full.ncu-rep --page raw --csv > $OUT/enc_attn_raw.csv 2>/dev/null
python - <<PY
import csv
rows = list(csv.reader(open("$OUT/enc_attn_raw.csv")))
hdr, units, r = rows[0], rows[1], rows[2]
keys = ["gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread"]
for i, h in enumerate(hdr):
    if h in keys or "pipe_xu" in h or "issue_active" in h or ("warp_issue_stalled" in h and "pct" in h) or "average_warp" in h and "stall" in h:
        print(h, "=", r[i], units[i])
PY
