#!/bin/bash
# final mini capture: default bench line, per-layer chained timings, ncu launch list
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r1h_bench_default.log 2>&1; echo "bench rc=$?"; tail -c 400 gpurun_out/r1h_bench_default.log
for m in dccrn dcunet20 frcrn; do timeout 300 python scripts/chain_layers.py --model $m; done > gpurun_out/r1h_chain_layers.txt 2>&1
grep TOTAL gpurun_out/r1h_chain_layers.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/r1h_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r1h_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/r1h_launches.csv
for m in dcunet20 frcrn; do timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r1h_bench_$m.log 2>&1; tail -1 gpurun_out/r1h_bench_$m.log | cut -c1-200; done
