#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload c5 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 rc=$?"; tail -3 gpurun_out/bench_c5_n$N.err | cut -c1-300; cat gpurun_out/bench_c5_n$N.json | cut -c1-1500
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-c3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "c2 rc=$?"; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
python - <<PY
import json
d = json.load(open("gpurun_out/bench_n$N.json"))
print(d["value"], d["us_per_rollout_step"], json.dumps(d.get("c4_cma_population"))[:900])
PY
