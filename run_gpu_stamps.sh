#!/bin/bash
# debug aid: rebuild with phase stamps, print them for the C3 and C2 forward
mkdir -p gpurun_out
make -s -C gnn_manip_b200/csrc clean all EXTRA="-DGM_STAMPS $GM_EXP" > gpurun_out/stamps_build.log 2>&1
python tools_stamps.py 100000 > gpurun_out/stamps_c3.txt 2>&1; tail -6 gpurun_out/stamps_c3.txt
python tools_stamps.py 2000 > gpurun_out/stamps_c2.txt 2>&1; tail -6 gpurun_out/stamps_c2.txt
