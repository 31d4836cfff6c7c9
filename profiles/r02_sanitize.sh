#!/bin/bash
# profiles/r02_sanitize.sh memcheck|racecheck|synccheck|initcheck  (ONE tool per gpurun call, B200_PROFILING.md)
# Note: late in round 2 the pool closed compute-sanitizer ("runs under it have left GPUs needing a reset"); the plain run of
# profiles/sanitize_case.py (every product kernel incl. one training step per env) still serves as the small-case smoke.
tool=${1:-memcheck}
mkdir -p gpurun_out
python profiles/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python profiles/sanitize_case.py > gpurun_out/sanitize_$tool.log 2>&1
echo "compute-sanitizer $tool exit $?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|sanitize case done|ok " gpurun_out/sanitize_$tool.log | tail -20
