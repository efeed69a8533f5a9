#!/bin/bash
# compute-sanitizer over the scoring path on small parity cases: ONE tool per gpurun call
# (B200_PROFILING.md), on a program that has just exited 0 without the tool.
#   gpurun --timeout 900 -- 'bash tools/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
CMD="python tests/parity_worker.py deep"
$CMD > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 800 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 3 $CMD > gpurun_out/sanitize_${TOOL}.log 2>&1
rc=$?
echo "compute-sanitizer --tool $TOOL rc=$rc"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|PARITY_OK|hazard" gpurun_out/sanitize_${TOOL}.log | head -10
exit $rc
