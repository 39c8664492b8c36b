#!/bin/bash
# re-entry check: full GPU parity suite, smoke, default bench and the two other model benches
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r25_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r25_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r25_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r25_smoke.log
for m in dccrn dcunet20 frcrn; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r25_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -c 1500 gpurun_out/r25_bench_$m.log | tail -1 | cut -c1-600
done
