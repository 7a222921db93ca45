#!/bin/bash
# whole GPU suite, then the quick C2 / C3 / C5 numbers
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 120 python bench.py --only-c3 2>/dev/null | cut -c1-220
timeout 200 python bench.py --no-c3 --no-c4 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
timeout 300 python bench.py --workload c5 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_q.json"))
print("C2 us/step", round(d["us_per_rollout_step"], 1), {k: round(v, 1) for k, v in d["kernel_us_c2"].items()})
d = json.load(open("gpurun_out/bench_c5.json"))
print("C5", round(d["value"], 1), "graphs/s", round(d["ms_per_step"], 2), "ms", d["split_ms"])
PY
