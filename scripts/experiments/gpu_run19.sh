#!/bin/bash
mkdir -p gpurun_out
for m in dcunet20 frcrn; do
  timeout 900 python bench.py --model $m --steps 5 --warmup 3 --batch 64 > gpurun_out/r22_bench_$m.log 2>&1
  tail -1 gpurun_out/r22_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['metric'], d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['achieved'], d['stages'], d['cpu_baseline']['value'])" || tail -5 gpurun_out/r22_bench_$m.log
done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r22_bench_dccrn.log 2>&1; tail -1 gpurun_out/r22_bench_dccrn.log | cut -c1-200
