#!/bin/bash
# whole GPU suite (no -x: list every failure), then optionally extra commands
set -u
TAG=${1:-tests}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout ${TEST_TIMEOUT:-1800} python -m pytest tests -m gpu -q --timeout 900 ${PYTEST_ARGS:-} > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" $OUT/pytest_gpu.log | head -40
