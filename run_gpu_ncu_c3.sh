#!/bin/bash
# ncu --set full of one kernel class on the C3 (100k particles) forward; usage: run_gpu_ncu_c3.sh <kernel regex> <skip> <out name>
mkdir -p gpurun_out
K=${1:-mp_edge}; S=${2:-12}; O=${3:-prof_$K}
C3_CMD="python bench.py --only-c3"
$C3_CMD > gpurun_out/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -f -o gpurun_out/$O $C3_CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_c3.log; tail -3 gpurun_out/ncu_full.log
