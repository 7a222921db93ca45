#!/bin/bash
mkdir -p gpurun_out
GNN_MANIP_B200_LIB=$PWD/gnn_manip_b200/libgnn_manip_b200_stamps.so timeout 300 python tools_stamps.py 2000 > gpurun_out/stamps_c2.txt 2>&1; tail -9 gpurun_out/stamps_c2.txt
GNN_MANIP_B200_LIB=$PWD/gnn_manip_b200/libgnn_manip_b200_stamps.so timeout 300 python tools_stamps.py 100000 > gpurun_out/stamps_c3.txt 2>&1; tail -9 gpurun_out/stamps_c3.txt
