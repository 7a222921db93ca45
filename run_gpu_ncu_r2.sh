#!/bin/bash
# round-2 ncu evidence: launch list of a short C2 rollout, then --set full of neighbour_kernel and mp_node16 on the C3 step
mkdir -p gpurun_out
NCU_CMD="python bench.py --steps 1 --warmup 3 --rollout-steps 4 --no-cpu-baseline --no-c3 --no-c4"
$NCU_CMD > gpurun_out/plain_small.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 1500 --csv --log-file gpurun_out/launches_r02.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
C3="python bench.py --only-c3-step"
$C3 > gpurun_out/plain_c3step.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:neighbour_kernel -s 1 -c 1 -f -o gpurun_out/prof_neighbour_c3 $C3 > gpurun_out/ncu_nb.log 2>&1
echo "neighbour rc=$?"; tail -2 gpurun_out/ncu_nb.log
ncu --set full --clock-control none --import-source on -k regex:mp_node16 -s 12 -c 1 -f -o gpurun_out/prof_mp_node16_c3 $C3 > gpurun_out/ncu_node.log 2>&1
echo "node16 rc=$?"; tail -2 gpurun_out/ncu_node.log
cat gpurun_out/plain_c3step.log | tail -2
