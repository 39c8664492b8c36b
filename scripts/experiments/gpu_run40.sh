#!/bin/bash
mkdir -p gpurun_out
SE_LSTM_COLS=3 timeout 900 python -m pytest tests -m gpu -x -q -k "lstm" > gpurun_out/r44_tests_lstm3.log 2>&1
echo "lstm3 tests rc=$?"; tail -5 gpurun_out/r44_tests_lstm3.log | cut -c1-220
for cols in 2 3 4 5; do
  SE_LSTM_COLS=$cols timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r44_bench_cols$cols.log 2>&1
  echo "bench cols=$cols rc=$?"; tail -1 gpurun_out/r44_bench_cols$cols.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['stages'].items() if not isinstance(v,dict)})" || tail -5 gpurun_out/r44_bench_cols$cols.log
done
