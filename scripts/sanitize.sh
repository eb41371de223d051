#!/usr/bin/env bash
# Race / memory checking for the native kernels (the reference ships none).
#
#   scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest args...]
#
# 1. compute-sanitizer over the single-GPU kernel tests (one tool per invocation; needs a GPU box where the tool is
#    allowed - on the shared B200 pool compute-sanitizer is closed, use steps 2 and 3 there);
# 2. DSB200_SYMM_VERIFY=1: every fused NVLink reduce-scatter inside the ZeRO optimizer is checked against the NCCL
#    collective of the same data (catches flag-protocol ordering bugs as numerical mismatches);
# 3. the loop-back tests of the collective GEMM (tests/test_gemm_comm_gpu.py) and the signal-protocol tests in
#    tests/test_symm_multigpu.py (buffer reuse across epochs, DONE waits).
set -euo pipefail
tool="${1:-memcheck}"
shift || true
cd "$(dirname "$0")/.."
if command -v compute-sanitizer >/dev/null 2>&1 && [ "${DSB200_NO_SANITIZER:-0}" != "1" ]; then
    compute-sanitizer --tool "$tool" --error-exitcode 1 \
        python -m pytest tests/test_kernels_gpu.py tests/test_token_ops_gpu.py tests/test_gemm_gpu.py -x -q -m gpu "$@" || exit 1
fi
DSB200_SYMM_VERIFY=1 python -m pytest tests/test_symm_multigpu.py tests/test_gemm_comm_gpu.py -x -q -m gpu "$@"
