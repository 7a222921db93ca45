#!/bin/bash
# usage: ./run_gpu_quicktest.sh <pytest args...>   (log -> gpurun_out/quicktest.log)
mkdir -p gpurun_out
timeout 900 python -m pytest "$@" > gpurun_out/quicktest.log 2>&1
echo "pytest rc=$?"; tail -n 60 gpurun_out/quicktest.log | cut -c1-300
