source tools/gpu_trace.sh ${TAG:-trace4}
timeout 600 python -m pytest tests/test_gpu_golden.py -x -q -k "lanes or batch" 2>&1 | tail -5
TAILN=4
run g2 WB200_LANES=1 WB200_GROUPS=2
run g3 WB200_LANES=1 WB200_GROUPS=3
run g4 WB200_LANES=1 WB200_GROUPS=4
run g3_l2 WB200_LANES=2 WB200_GROUPS=3 WB200_PAIR_MIN=40
BATCH=384 run g3_b384 WB200_LANES=1 WB200_GROUPS=3
