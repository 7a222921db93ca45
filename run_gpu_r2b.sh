#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 5 gpurun_out/bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
for k in ("value", "us_per_rollout_step", "e2e", "c4_cma_population", "roofline", "kernel_us_c2", "kernel_us_c3", "phase_us_c3", "c3_rollout", "cpu_baseline"):
    print(k, json.dumps(d.get(k))[:700])
for k, v in (d.get("rooflines_c3") or {}).items():
    print(k, round(v["us"], 1), "us", round(v["achieved"], 1), "GB/s frac", round(v["frac"], 4))
print(d["config"])
PY
