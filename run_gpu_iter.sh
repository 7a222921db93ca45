#!/bin/bash
# iteration loop on the GPU box: parity suite, quick bench line, then the phase stamps of the edge kernel
mkdir -p gpurun_out
if [ -z "$GM_SKIP_TESTS" ]; then timeout 600 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log; fi
timeout 600 python bench.py --no-cpu-baseline "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_quick.json"))
    print("C2 us/step", round(d["us_per_rollout_step"], 1), "value", round(d["value"]), "kernel_us_c2", {k: round(v, 1) for k, v in d["kernel_us_c2"].items()})
    print("C3 kernel_us", {k: round(v, 1) for k, v in d["kernel_us_c3"].items()}, "c3 pps", round(d["c3_model_particle_steps_per_s"]))
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 5 gpurun_out/bench_quick.err
make -s -C gnn_manip_b200/csrc clean all EXTRA=-DGM_STAMPS > gpurun_out/stamps_build.log 2>&1
timeout 300 python tools_stamps.py 100000 > gpurun_out/stamps_c3.txt 2>&1; tail -6 gpurun_out/stamps_c3.txt
if [ -n "$GM_EXP" ]; then
  make -s -C gnn_manip_b200/csrc clean all EXTRA="-DGM_STAMPS $GM_EXP" > gpurun_out/stamps_build2.log 2>&1
  echo "== experiment build: $GM_EXP"
  timeout 300 python tools_stamps.py 100000 > gpurun_out/stamps_c3_exp.txt 2>&1; tail -6 gpurun_out/stamps_c3_exp.txt | head -3
fi
