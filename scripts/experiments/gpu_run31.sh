#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "lstm or dccrn or frcrn" > gpurun_out/r35_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r35_tests.log
for rows in 0 -1 4 8; do
  SE_LSTM_ROWS=$rows timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r35_bench_rows$rows.log 2>&1
  echo "rows=$rows rc=$?"; tail -1 gpurun_out/r35_bench_rows$rows.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], 'lstm', d['stages']['lstm_recurrence_ms'])"
done
timeout 600 python scripts/graph_bench.py 2>&1 | tail -5
