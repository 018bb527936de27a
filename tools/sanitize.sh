#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests (SURVEY §5 race detection).  One tool per invocation and per gpurun call:
#   gpurun --timeout 900 -- 'python -m pytest tests/test_segment_ops_gpu.py -x -q && tools/sanitize.sh memcheck tests/test_segment_ops_gpu.py'
# Tools: memcheck | racecheck | synccheck | initcheck.  Run the plain test first (same call, &&) so a sanitizer failure is
# not confused with an ordinary test failure.
set -euo pipefail
tool="${1:-memcheck}"; shift || true
targets=("${@:-tests/test_segment_ops_gpu.py}")
mkdir -p gpurun_out
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
exec compute-sanitizer --tool "$tool" --error-exitcode 99 --print-limit 20 --launch-timeout 0 \
  --log-file "gpurun_out/sanitizer_${tool}.log" \
  python -m pytest "${targets[@]}" -x -q -m gpu -p no:cacheprovider
