#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_kernels.py tests/test_gpu_formats.py -q -m gpu -s > gpurun_out/t_train.log 2>&1; echo "pytest rc=$?"; grep "^n=\|^\.n=\|Adam 3\|passed\|failed" gpurun_out/t_train.log | cut -c1-400
timeout 900 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"; tail -5 gpurun_out/bench_c5.err; cat gpurun_out/bench_c5.json
