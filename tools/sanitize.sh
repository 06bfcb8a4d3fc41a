#!/bin/bash
# compute-sanitizer over the native kernels.  ONE tool per invocation (B200_PROFILING.md: several tools in one gpurun call
# have left GPUs unusable):   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck
# Run it on the GPU box:      gpurun -- 'tools/sanitize.sh racecheck'
# The summary goes to gpurun_out/sanitize_<tool>.log; copy the verdict lines into profiles/sanitizer.md.
set -u
TOOL="${1:-memcheck}"
mkdir -p gpurun_out
python tools/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "target fails without the sanitizer"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool "$TOOL" --print-limit 20 --error-exitcode 3 python tools/sanitize_target.py > "gpurun_out/sanitize_${TOOL}.log" 2>&1
rc=$?
echo "compute-sanitizer --tool $TOOL exit code $rc"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|sanitize target ok" "gpurun_out/sanitize_${TOOL}.log" | tail -15
exit 0
