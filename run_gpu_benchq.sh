#!/bin/bash
# quick C2 numbers: bench without C3 / C4 / CPU legs
mkdir -p gpurun_out
timeout 600 python bench.py --no-c3 --no-c4 --no-cpu-baseline "$@" > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_q.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_q.json"))
print("C2 us/step", round(d["us_per_rollout_step"], 1), "value", round(d["value"]), "kernel_us_c2", {k: round(v, 1) for k, v in d["kernel_us_c2"].items()})
PY
