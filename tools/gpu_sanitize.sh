#!/bin/bash
# compute-sanitizer over the small bf16 block cases; ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun --timeout 900 -- 'bash tools/gpu_sanitize.sh memcheck'      then      ... racecheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 120 python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 700 compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitize_$TOOL.log python tools/sanitize_case.py > gpurun_out/sanitize_${TOOL}_stdout.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/sanitize_${TOOL}_stdout.log; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" gpurun_out/sanitize_$TOOL.log | sort | uniq -c | head -20
