#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -n 4 gpurun_out/noalloc_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log | cut -c1-300
