#!/bin/bash
# launch list of ONE training step (C5): plain run first, then ncu per-launch durations
mkdir -p gpurun_out
CMD="python tools_train_launches.py"
timeout 300 $CMD > gpurun_out/plain_c5.log 2>&1 && timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_c5.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_c5.log | cut -c1-300
python profiles/summarise_launches.py gpurun_out/launches_c5.csv | head -16
