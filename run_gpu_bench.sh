#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
NCU_CMD="python bench.py --steps 1 --warmup 3 --rollout-steps 4 --no-cpu-baseline --no-c3 --no-c4"
$NCU_CMD > gpurun_out/plain_small.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
C3_CMD="python bench.py --only-c3"
$C3_CMD > gpurun_out/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mp_edge -s 12 -c 2 -o gpurun_out/prof_mp_edge_c3 $C3_CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
