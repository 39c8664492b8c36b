#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "lstm or dccrn or frcrn" > gpurun_out/r24_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r24_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r24_bench.log 2>&1
tail -1 gpurun_out/r24_bench.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['metric'], d['value'], d['ms_per_step'], d['e2e']['value'], d['stages'])" || tail -5 gpurun_out/r24_bench.log
