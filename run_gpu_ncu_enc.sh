#!/bin/bash
mkdir -p gpurun_out
C3="python bench.py --only-c3"
timeout 120 $C3 > gpurun_out/plain_c3.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:enc_edge_pair -s 1 -c 1 -f -o gpurun_out/prof_enc_edge $C3 > gpurun_out/ncu_enc.log 2>&1
echo "enc rc=$?"
