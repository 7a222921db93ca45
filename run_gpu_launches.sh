#!/bin/bash
# ncu launch list (device time per launch) of a short C2 rollout; caches left warm
mkdir -p gpurun_out
NCU_CMD="python bench.py --steps 1 --warmup 3 --rollout-steps 4 --no-cpu-baseline --no-c3"
$NCU_CMD > gpurun_out/plain_small.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 1500 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
