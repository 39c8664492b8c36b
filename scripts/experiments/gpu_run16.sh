#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "lstm or dccrn or cconv" > gpurun_out/r17_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r17_tests.log
SE_GEMM_PROF=1 SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/r17_prof.log 2>&1
grep "se_cconv_gemm prof" gpurun_out/r17_prof.log | tail -15
SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r17_bench.log 2>&1; tail -1 gpurun_out/r17_bench.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages'])"
