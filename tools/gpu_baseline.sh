#!/bin/bash
# One GPU-box session: parity tests, the default bench line, then ncu evidence (launch list + full capture of the top kernels).
# Usage (from the repo root, through gpurun):  bash tools/gpu_baseline.sh <tag>
set -u
TAG=${1:-r1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" | tee -a $OUT/smoke.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; tail -c 3000 $OUT/bench.json
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$BCMD > $OUT/bench_short.json 2> $OUT/bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 4000 -c 3000 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"gemm_f16_tcgen05|cross_attn_kernel" -s 3000 -c 8 -o $OUT/prof_top $BCMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
