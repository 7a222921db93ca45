#!/bin/bash
# debug aid: rebuild with phase stamps, print them for the C3 and C2 forward, rebuild the normal library
mkdir -p gpurun_out
make -s -C gnn_manip_b200/csrc clean all EXTRA=-DGM_STAMPS > gpurun_out/stamps_build.log 2>&1
python tools_stamps.py 100000 > gpurun_out/stamps_c3.txt 2>&1; cat gpurun_out/stamps_c3.txt | tail -8
python tools_stamps.py 2000 > gpurun_out/stamps_c2.txt 2>&1; cat gpurun_out/stamps_c2.txt | tail -8
