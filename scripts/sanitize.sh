#!/bin/bash
# compute-sanitizer over the kernel tests.  NOTE: the tool is closed on the shared B200 pool this repo was built on
# (gpurun answers rc=86); use it where it is allowed.  ONE tool per gpurun call (B200_PROFILING.md); the plain run goes first.
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck tests/test_resnet_kernels.py'
TOOL=${1:-memcheck}; shift
TARGETS=${@:-tests/test_resnet_kernels.py tests/test_flat_ops.py}
mkdir -p gpurun_out
timeout 300 python -m pytest $TARGETS -m gpu -x -q > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 800 compute-sanitizer --tool "$TOOL" --error-exitcode 1 --print-limit 20 python -m pytest $TARGETS -m gpu -x -q > gpurun_out/sanitize_$TOOL.log 2>&1
rc=$?
echo "compute-sanitizer --tool $TOOL rc=$rc"
grep -E "ERROR SUMMARY|passed|failed|Invalid|Race|hazard" gpurun_out/sanitize_$TOOL.log | tail -8
exit $rc
