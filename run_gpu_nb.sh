#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_regimes.py -q -m gpu -x > gpurun_out/nb_tests.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/nb_tests.log | cut -c1-300
timeout 200 python bench.py --only-c3-step 2>/dev/null | tail -2 | cut -c1-900
timeout 200 python bench.py --no-c3 --no-c4 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_q.json"))
print("C2 us/step", round(d["us_per_rollout_step"], 1), {k: round(v, 1) for k, v in d["kernel_us_c2"].items()})
PY
