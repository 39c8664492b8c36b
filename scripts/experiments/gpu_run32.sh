#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r36_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r36_tests.log
for m in dccrn dcunet20 frcrn; do timeout 300 python scripts/chain_layers.py --model $m 2>&1 | grep -E "first|TOTAL"; done
for m in dccrn dcunet20 frcrn; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r36_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -1 gpurun_out/r36_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
done
