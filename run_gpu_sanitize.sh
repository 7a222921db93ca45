#!/bin/bash
mkdir -p gpurun_out
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest "$@" > gpurun_out/sanitize.log 2>&1; echo "rc=$?"
grep -n "Invalid\|at 0x\|by thread\|Address\|ERROR SUMMARY\|passed\|failed" gpurun_out/sanitize.log | head -30
