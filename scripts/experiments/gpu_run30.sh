#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r34_tests.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r34_tests.log
for m in dccrn dcunet20 frcrn; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r34_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -1 gpurun_out/r34_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:(round(v['ms'],3), round(v.get('frac_hbm',0) or 0,3)) for k,v in d['stages'].items() if isinstance(v,dict) and 'ms' in v})"
done
