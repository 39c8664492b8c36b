#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r48_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r48_tests.log | cut -c1-220
for m in dccrn dcunet20 frcrn; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r48_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -1 gpurun_out/r48_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks']['sm_mhz'], d['cpu_baseline']['value'], {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['stages'].items() if not isinstance(v,dict)})" || tail -5 gpurun_out/r48_bench_$m.log
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r48_bench_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/r48_bench_ref.log | cut -c1-400
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
