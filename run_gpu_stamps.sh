#!/bin/bash
# debug aid: rebuild with phase stamps, print them for the C3 and C2 forward, and for a graph rollout step
mkdir -p gpurun_out
make -s -C gnn_manip_b200/csrc clean all EXTRA="-DGM_STAMPS $GM_EXP" > gpurun_out/stamps_build.log 2>&1
python tools_stamps.py 100000 > gpurun_out/stamps_c3.txt 2>&1; tail -7 gpurun_out/stamps_c3.txt
python tools_stamps_graph.py 2000 graph 2>&1 | tail -1
python tools_stamps_graph.py 2000 nograph 2>&1 | tail -1
