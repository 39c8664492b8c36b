#!/bin/bash
mkdir -p gpurun_out
for ck in 256 128 64; do
  SE_LINEAR_CHUNK=$ck timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r49_bench_ck$ck.log 2>&1
  echo "bench chunk=$ck rc=$?"; tail -1 gpurun_out/r49_bench_ck$ck.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks']['sm_mhz'], {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['stages'].items() if not isinstance(v,dict)})" || tail -5 gpurun_out/r49_bench_ck$ck.log
done
SE_LINEAR_CHUNK=128 timeout 600 python -m pytest tests -m gpu -x -q -k "lstm or linear" 2>&1 | tail -2
