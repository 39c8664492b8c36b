#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "lstm or linear or dccrn or frcrn" > gpurun_out/r45_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r45_tests.log | cut -c1-220
for ns in 0 2 4 8; do
  SE_LINEAR_STREAMS=$ns timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r45_bench_ns$ns.log 2>&1
  echo "bench streams=$ns rc=$?"; tail -1 gpurun_out/r45_bench_ns$ns.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['stages'].items() if not isinstance(v,dict)})" || tail -5 gpurun_out/r45_bench_ns$ns.log
done
