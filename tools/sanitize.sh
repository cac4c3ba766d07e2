#!/bin/bash
# compute-sanitizer job on the smallest end-to-end case (one tool per gpurun call, see
# /opt/skills/guides/B200_PROFILING.md).  Usage: bash tools/sanitize.sh memcheck|racecheck|synccheck|initcheck
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain smoke failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitize_$TOOL.log python __graft_entry__.py smoke > gpurun_out/sanitize_${TOOL}_stdout.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"; tail -4 gpurun_out/sanitize_$TOOL.log
