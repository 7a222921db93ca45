#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train_kernels.py tests/test_gpu_train.py tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/train_tests.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/train_tests.log | cut -c1-300
timeout 300 python bench.py --workload c5 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_c5.json"))
print(d["value"], d["ms_per_step"], d["split_ms"], d["gpu_launches"], d["loss"])
PY
