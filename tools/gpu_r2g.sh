#!/bin/bash
# sanitizer (one tool per call: $1), bounded
mkdir -p gpurun_out
TOOL=${1:-memcheck}
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/sanitize_case.py > gpurun_out/sanitizer_$TOOL.log 2>&1; echo "sanitizer $TOOL rc=$?"
grep -E "ok$|ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" gpurun_out/sanitizer_$TOOL.log | head -30
