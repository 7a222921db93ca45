#!/bin/bash
# first-contact GPU run: each group in its own process so a CUDA fault in one does not poison the rest
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for grp in "selftest_gemm" "graph" "features or next_position" "blocks or scatter" "forward and 1-" "forward and 2-" "forward and not 1- and not 2-" "generic" "rollout"; do
  name=$(echo "$grp" | tr ' ' '_' )
  timeout 240 python -m pytest tests -q -m gpu -k "$grp" -x -s > "gpurun_out/t_${name}.log" 2>&1
  echo "== $grp -> exit $?" | tee -a gpurun_out/summary.txt
  tail -n 4 "gpurun_out/t_${name}.log"
done
