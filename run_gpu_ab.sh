#!/bin/bash
# parity of the forward / rollout tests, then quick C2 numbers for the default build and with the env override in $1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_regimes.py -q -m gpu -x > gpurun_out/ab_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ab_tests.log
./run_gpu_benchq.sh
if [ -n "$1" ]; then env $1 ./run_gpu_benchq.sh; fi
