#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r23_tests.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r23_tests.log
for m in dccrn dcunet20; do
  timeout 900 python bench.py --model $m --steps 5 --warmup 3 --batch 64 > gpurun_out/r23_bench_$m.log 2>&1
  tail -1 gpurun_out/r23_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['metric'], d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['achieved'], d['stages']['cconv'])" || tail -5 gpurun_out/r23_bench_$m.log
done
