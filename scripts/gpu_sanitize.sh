#!/bin/bash
# One compute-sanitizer tool per gpurun call (B200_PROFILING.md), on small parity tests.
TOOL=${1:-memcheck}
python -m pytest tests/test_golden.py -m gpu -q -x > gpurun_out/plain.log 2>&1 && \
compute-sanitizer --tool $TOOL --error-exitcode 9 python -m pytest tests/test_golden.py -m gpu -q -x > gpurun_out/sanitize_$TOOL.log 2>&1
echo "sanitizer $TOOL exit=$?"; tail -5 gpurun_out/sanitize_$TOOL.log
