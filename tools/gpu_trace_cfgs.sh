source tools/gpu_trace.sh ${TAG:-trace5}
TAILN=4
run l1 WB200_LANES=1
BATCH=384 run l4_b384 WB200_LANES=4
