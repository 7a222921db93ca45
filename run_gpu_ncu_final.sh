#!/bin/bash
# round-2 final ncu evidence: C2 launch list, --set full of mp_edge16 (C3 forward), neighbour_kernel (C3 step), tr_gemm (C5 step)
mkdir -p gpurun_out
NCU_CMD="python bench.py --steps 1 --warmup 3 --rollout-steps 4 --no-cpu-baseline --no-c3 --no-c4"
timeout 200 $NCU_CMD > gpurun_out/plain_small.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 1500 --csv --log-file gpurun_out/launches_r02_final.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
C3="python bench.py --only-c3"
timeout 120 $C3 > gpurun_out/plain_c3.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:mp_edge16 -s 4 -c 1 -f -o gpurun_out/prof_mp_edge16_final $C3 > gpurun_out/ncu_edge.log 2>&1
echo "edge16 rc=$?"
C3S="python bench.py --only-c3-step"
timeout 120 $C3S > gpurun_out/plain_c3step.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:neighbour_kernel -s 1 -c 1 -f -o gpurun_out/prof_neighbour_final $C3S > gpurun_out/ncu_nb.log 2>&1
echo "neighbour rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mp_node16 -s 4 -c 1 -f -o gpurun_out/prof_mp_node16_final $C3 > gpurun_out/ncu_node.log 2>&1
echo "node16 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:enc_edge_pair -s 1 -c 1 -f -o gpurun_out/prof_enc_edge_final $C3 > gpurun_out/ncu_enc.log 2>&1
echo "enc_edge rc=$?"
TR="python tools_train_launches.py"
timeout 200 $TR > gpurun_out/plain_c5.log 2>&1 && timeout 400 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:tr_gemm --launch-skip 7 -c 1 -f -o gpurun_out/prof_tr_gemm_final $TR > gpurun_out/ncu_gemm.log 2>&1
echo "tr_gemm rc=$?"
timeout 400 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches_c5_final.csv $TR > gpurun_out/ncu_c5.log 2>&1
echo "c5 launch list rc=$?"
