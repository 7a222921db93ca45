#!/bin/bash
# full GPU suite, smoke, both bench arms, training bench -> gpurun_out/
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench.err
timeout 600 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; cut -c1-600 gpurun_out/bench_c5.json
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
for k in ("value", "us_per_rollout_step", "e2e", "c4_cma_population", "roofline", "kernel_us_c2", "kernel_us_c3", "phase_us_c3", "c3_rollout", "cpu_baseline", "gpu_launches"):
    print(k, json.dumps(d.get(k))[:700])
PY
