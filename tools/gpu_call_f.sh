#!/bin/bash
# one sanitizer tool per call (B200_PROFILING.md): $1 = memcheck | racecheck
mkdir -p gpurun_out
timeout 300 python tools/sanitize_target.py > gpurun_out/f_plain_$1.log 2>&1 || { tail -20 gpurun_out/f_plain_$1.log; exit 1; }
timeout 2400 compute-sanitizer --tool $1 --print-limit 50 python tools/sanitize_target.py > gpurun_out/sanitizer_$1.log 2>&1
echo "sanitizer rc=$?" >> gpurun_out/sanitizer_$1.log
tail -25 gpurun_out/sanitizer_$1.log
