#!/bin/bash
mkdir -p gpurun_out
T="tests/test_gpu_traj.py -q -m gpu -x -k cma_objective_population"
for v in "A=1" "GM_NODE_KERNEL=8" "GM_GRAPH_TINY=0" "CUDA_LAUNCH_BLOCKING=1"; do
  env $v timeout 300 python -m pytest $T > gpurun_out/bisect_$v.log 2>&1; echo "$v rc=$?"; grep -n "passed\|failed\|Error" gpurun_out/bisect_$v.log | head -3
done
grep -n "failed (status\|gm_\|\.py:[0-9]*: " "gpurun_out/bisect_CUDA_LAUNCH_BLOCKING=1.log" | head
