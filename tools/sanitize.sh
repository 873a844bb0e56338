#!/bin/bash
# tools/sanitize.sh TOOL   (memcheck | racecheck | initcheck | synccheck) — ONE tool per gpurun call (B200_PROFILING.md).
# Runs tools/sanitize_case.py (every kernel family once, small sizes, results checked against the oracle) under
# compute-sanitizer; the log goes to gpurun_out/r02_sanitizer_TOOL.log and its summary lines are printed.
TOOL=${1:-memcheck}
python tools/sanitize_case.py > gpurun_out/r02_sanitizer_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/r02_sanitizer_plain.log; exit 1; }
EXTRA=""
[ "$TOOL" = "racecheck" ] && EXTRA="--racecheck-report all"
[ "$TOOL" = "memcheck" ] && EXTRA="--leak-check no"
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --error-exitcode 9 --print-limit 40 python tools/sanitize_case.py > gpurun_out/r02_sanitizer_$TOOL.log 2>&1
echo "compute-sanitizer --tool $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|== oracle|sanitize_case ok|hazard|Invalid|Uninitialized|Barrier error" gpurun_out/r02_sanitizer_$TOOL.log | sort | uniq -c | sort -rn | head -30
