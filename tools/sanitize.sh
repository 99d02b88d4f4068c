#!/bin/bash
# compute-sanitizer over the single-GPU kernel self-check.  ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun -- 'bash tools/sanitize.sh memcheck'        (or racecheck / synccheck / initcheck)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
DLBB_SANITIZE=1 compute-sanitizer --tool $TOOL --print-limit 20 python tools/gpu_sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL exit code $?"; tail -15 gpurun_out/sanitize_$TOOL.log
