#!/bin/bash
# round-2 first contact: the new regime tests (each group in its own process), then the whole suite time, then a bench line
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_regimes.py -q -m gpu -s --durations=0 > gpurun_out/t_regimes.log 2>&1
echo "regimes rc=$?"; tail -n 40 gpurun_out/t_regimes.log | cut -c1-400
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu > gpurun_out/t_parity.log 2>&1
echo "parity rc=$?"; tail -n 3 gpurun_out/t_parity.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err; echo "bench rc=$?"
cat gpurun_out/bench_r2a.json | cut -c1-1500
