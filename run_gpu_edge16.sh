#!/bin/bash
# A/B of the edge kernels: C3 forward profile first (short timeout: a hang must not eat the budget), parity tests, C2 quick bench
mkdir -p gpurun_out
for k in 16 8; do
  GM_EDGE_KERNEL=$k timeout 120 python bench.py --only-c3 2>gpurun_out/only_c3_$k.err | cut -c1-400; echo "only-c3 $k rc=${PIPESTATUS[0]}"
done
tail -3 gpurun_out/only_c3_16.err | cut -c1-300
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_regimes.py -q -m gpu -x > gpurun_out/edge16_tests.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/edge16_tests.log | cut -c1-300
for k in 16 8; do
  GM_EDGE_KERNEL=$k timeout 200 python bench.py --no-c3 --no-c4 --no-cpu-baseline > gpurun_out/bench_q$k.json 2> gpurun_out/bench_q$k.err; echo "bench rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_q$k.json"))
    print("edge kernel $k: C2 us/step", round(d["us_per_rollout_step"], 1), "kernel_us_c2", {k: round(v, 1) for k, v in d["kernel_us_c2"].items()})
except Exception as e:
    print("no bench line", e)
PY
done
