#!/bin/bash
# usage: tools/gpu_sanitize.sh memcheck|racecheck   (one tool per gpurun call)
TOOL=$1
mkdir -p gpurun_out
python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/sanitize_case.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "rc=$?"; tail -8 gpurun_out/sanitize_$TOOL.log
