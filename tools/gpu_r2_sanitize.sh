#!/bin/bash
# compute-sanitizer pass: bash tools/gpu_r2_sanitize.sh memcheck|racecheck|synccheck|initcheck [cases...]
tool=${1:-memcheck}; shift
mkdir -p gpurun_out
timeout 1700 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py "$@" > gpurun_out/r2_sanitizer_$tool.log 2>&1
echo "rc=$?" >> gpurun_out/r2_sanitizer_$tool.log
grep -E "sanitize\]|ERROR SUMMARY|RACECHECK SUMMARY|Error|rc=" gpurun_out/r2_sanitizer_$tool.log | tail -30
