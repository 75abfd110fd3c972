#!/bin/sh
# Race / memory checking of the hand-written kernels on a GPU box (one tool per gpurun call:
# /opt/skills/guides/B200_PROFILING.md).  Usage: scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck
set -e
cd "$(dirname "$0")/.."
TOOL="${1:-memcheck}"
exec compute-sanitizer --tool "$TOOL" --error-exitcode 9 --launch-timeout 0 \
    python -m pytest tests/test_kernels_gpu.py tests/test_gemm_gpu.py -m gpu -x -q -k "not unbiased"
