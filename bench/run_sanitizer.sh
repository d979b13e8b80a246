#!/bin/bash
# compute-sanitizer pass over the hand-written kernels (one tool per gpurun call).
#   gpurun --timeout 900 -- 'bash bench/run_sanitizer.sh memcheck'
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 200 python bench/sanitize_smoke.py > gpurun_out/sanitize_plain.log 2>&1
rc=$?; echo "plain run exit=$rc $(tail -n 2 gpurun_out/sanitize_plain.log | cut -c1-300)"
[ $rc -ne 0 ] && exit $rc
timeout 800 compute-sanitizer --tool "$TOOL" --error-exitcode 9 --log-file gpurun_out/sanitizer_$TOOL.log \
  python bench/sanitize_smoke.py > gpurun_out/sanitize_$TOOL.out 2>&1
echo "compute-sanitizer $TOOL exit=$? $(tail -n 4 gpurun_out/sanitizer_$TOOL.log | cut -c1-300)"
