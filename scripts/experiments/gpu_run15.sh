#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "lstm or dccrn" > gpurun_out/r16_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r16_tests.log
SE_GEMM_PROF=1 SE_GEMM_DEBUG=4 SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/r16_prof.log 2>&1
grep -A1 "se_cconv_gemm: " gpurun_out/r16_prof.log | tail -36
SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r16_bench.log 2>&1; tail -1 gpurun_out/r16_bench.log
SE_BENCH_BATCH=16 timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r16_bench16.log 2>&1; tail -1 gpurun_out/r16_bench16.log | cut -c1-400
