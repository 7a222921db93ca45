#!/bin/bash
mkdir -p gpurun_out
GNN_MANIP_B200_LIB=$PWD/gnn_manip_b200/libgnn_manip_b200_trace.so timeout 300 python tools_trace.py 2000 1 > gpurun_out/trace_c2.txt 2>&1; head -48 gpurun_out/trace_c2.txt
