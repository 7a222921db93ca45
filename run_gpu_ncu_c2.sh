#!/bin/bash
# ncu --set full of mp_edge launches inside the C2 rollout graph (DRAM traffic of the C2 roofline entry)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --rollout-steps 4 --no-cpu-baseline --no-c3 --no-c4"
$CMD > gpurun_out/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mp_edge_kernel -s 60 -c 2 -f -o gpurun_out/prof_mp_edge_c2 $CMD > gpurun_out/ncu_c2.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_c2.log
