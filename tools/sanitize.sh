#!/bin/bash
# compute-sanitizer on small shapes (ONE tool per GPU-box visit, as the profiling guide requires):
#   tools/sanitize.sh racecheck | synccheck | memcheck      -> gpurun_out/sanitize_<tool>.log
set -u
TOOL=${1:-racecheck}
mkdir -p gpurun_out
timeout 120 python tools/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1200 compute-sanitizer --tool $TOOL --print-limit 40 --launch-timeout 0 python tools/sanitize_cases.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "sanitizer rc=$?"
grep -E "^ok|all cases|ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error" gpurun_out/sanitize_$TOOL.log | sort | uniq -c | sort -rn | head -30
