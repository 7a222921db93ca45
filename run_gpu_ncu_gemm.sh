#!/bin/bash
# ncu --set full of one edge-level tr_gemm launch of a training step (the 8th tr_gemm launch: layer 2 of MP step 0)
mkdir -p gpurun_out
CMD="python tools_train_launches.py"
timeout 300 $CMD > gpurun_out/plain_c5.log 2>&1 && timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:tr_gemm --launch-skip 7 -c 1 -f -o gpurun_out/prof_tr_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_gemm.log | cut -c1-300
