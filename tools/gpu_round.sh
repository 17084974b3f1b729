#!/bin/bash
# round-end style GPU session: smoke, full GPU test suite, default bench (with CPU baseline), then the ncu evidence for the same command:
# launch list (gpu__time_duration) and one --set full capture of the dominant decode kernel.
set -u
TAG=${1:-round}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/smoke.log
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"; tail -4 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; tail -c 5500 $OUT/bench.json; tail -3 $OUT/bench.err
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ln_reduce|cross_attn_step|gemm_wstream|gemm_wcluster|self_attn_step|sampler" -s 3000 -c 1500 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"; python tools/ncu_launches.py $OUT/launches.csv | head -20
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cross_attn_step -s 200 -c 2 -o $OUT/cross_attn_full $BCMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"; ls -la $OUT/*.ncu-rep 2>/dev/null
ncu -i $OUT/cross_attn_full.ncu-rep --page raw --csv > $OUT/cross_attn_full_raw.csv 2>/dev/null && python tools/ncu_raw.py $OUT/cross_attn_full_raw.csv | head -20
