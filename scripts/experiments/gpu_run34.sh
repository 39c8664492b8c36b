#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
  --log-file gpurun_out/r38_launches_frcrn.csv python bench.py --model frcrn --steps 2 --warmup 3 > gpurun_out/r38_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/r38_launches_frcrn.csv
