#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "ccbam or frcrn" > gpurun_out/r39_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r39_tests.log | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 --model frcrn > gpurun_out/r39_bench_frcrn.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/r39_bench_frcrn.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], 'ccbam', d['stages']['ccbam_ms'], 'cconv', d['stages']['cconv']['ms'])"
