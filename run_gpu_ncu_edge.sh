#!/bin/bash
# ncu --set full of the edge MP kernel (C3 forward profile); usage: ./run_gpu_ncu_edge.sh <variant 16|8> <out name>
mkdir -p gpurun_out
C3="python bench.py --only-c3"
k=${1:-16}; o=${2:-prof_mp_edge${k}_c3}
GM_EDGE_KERNEL=$k timeout 120 $C3 > gpurun_out/plain_c3_$k.log 2>&1 && GM_EDGE_KERNEL=$k timeout 400 ncu --set full --clock-control none --import-source on -k regex:mp_edge -s 4 -c 1 -f -o gpurun_out/$o $C3 > gpurun_out/ncu_edge$k.log 2>&1
echo "edge$k rc=$?"; tail -2 gpurun_out/ncu_edge$k.log | cut -c1-300
