#!/bin/bash
# quick GPU check: operand-path self tests first (bounded), then the parity suite, then the bench line (no ncu)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -q -m gpu -x -k selftest > gpurun_out/pytest_selftest.log 2>&1; echo "selftest rc=$?"; tail -n 15 gpurun_out/pytest_selftest.log
timeout 600 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_quick.json; tail -n 5 gpurun_out/bench_quick.err
