#!/bin/bash
# ONE compute-sanitizer tool per gpurun call (B200_PROFILING.md):  bash tools/sanitize.sh memcheck|racecheck|initcheck [tag]
set -u
TOOL=${1:?tool}
TAG=${2:-r02}
mkdir -p gpurun_out
timeout 300 python tools/sanitize_driver.py > gpurun_out/${TAG}_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool ${TOOL} --print-limit 40 --log-file gpurun_out/${TAG}_sanitize_${TOOL}.log \
    python tools/sanitize_driver.py > gpurun_out/${TAG}_sanitize_${TOOL}.out 2>&1
echo "exit $?"
tail -5 gpurun_out/${TAG}_sanitize_${TOOL}.out
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Internal Sanitizer Error" gpurun_out/${TAG}_sanitize_${TOOL}.log | tail -3
grep -cE "^=========.*(Invalid|Uninitialized|hazard|Race)" gpurun_out/${TAG}_sanitize_${TOOL}.log
