#!/bin/bash
# compute-sanitizer over the small end-to-end case (tools/sanitize_case.py).  ONE tool per gpurun call
# (B200_PROFILING.md: the four tools in one call left a GPU unusable): tools/sanitize.sh memcheck|racecheck|initcheck|synccheck [case]
# Graphs off (plain launches), programmatic dependent launch ON.  The report lands in gpurun_out/sanitize_<tool>.log.
set -u
TOOL=${1:-memcheck}
CASE=${2:-all}
mkdir -p gpurun_out
export F5B_NO_GRAPH=1
# the plain run first: a program that faults must not be handed to the sanitizer
timeout 600 python tools/sanitize_case.py "$CASE" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
EXTRA=""
[ "$TOOL" = "memcheck" ] && EXTRA="--leak-check no"
[ "$TOOL" = "racecheck" ] && EXTRA="--racecheck-report all"
timeout 2400 /usr/local/cuda/bin/compute-sanitizer --tool "$TOOL" $EXTRA --print-limit 50 --launch-timeout 600 \
  python tools/sanitize_case.py "$CASE" > "gpurun_out/sanitize_${TOOL}.log" 2>&1
echo "sanitizer rc=$?" | tee -a "gpurun_out/sanitize_${TOOL}.log"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|max-abs|done" "gpurun_out/sanitize_${TOOL}.log" | tail -20
