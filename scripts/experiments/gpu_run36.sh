#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "pipeline or smoke or dccrn" > gpurun_out/r40_tests.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r40_tests.log | cut -c1-200
for m in dccrn dcunet20 frcrn; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r40_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -1 gpurun_out/r40_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e'])" || tail -5 gpurun_out/r40_bench_$m.log
done
