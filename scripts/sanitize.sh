#!/usr/bin/env bash
# compute-sanitizer on the small case, ONE tool per call (default memcheck):  scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck]
cd "$(dirname "${BASH_SOURCE[0]}")/.."; mkdir -p gpurun_out
TOOL="${1:-memcheck}"
python scripts/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 600 compute-sanitizer --tool "$TOOL" --error-exitcode 66 python scripts/sanitize_case.py > gpurun_out/sanitize_$TOOL.log 2>&1
rc=$?; tail -6 gpurun_out/sanitize_$TOOL.log; echo "compute-sanitizer $TOOL exit code $rc"
