#!/bin/bash
# tensor-core STFT/iSTFT bring-up: parity, timing, then the whole GPU suite and a bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "stft" > gpurun_out/r14_stft_tests.log 2>&1
echo "stft tests rc=$?" ; tail -5 gpurun_out/r14_stft_tests.log
SE_GEMM_DEBUG=4 timeout 300 python scripts/stft_bench.py --batch 4 > gpurun_out/r14_plan.log 2>&1
grep -m8 "TH=" gpurun_out/r14_plan.log
timeout 300 python scripts/stft_bench.py > gpurun_out/r14_stft_bench.log 2>&1; cat gpurun_out/r14_stft_bench.log
timeout 300 python scripts/stft_bench.py --batch 16 >> gpurun_out/r14_stft_bench.log 2>&1; tail -2 gpurun_out/r14_stft_bench.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r14_tests.log 2>&1
echo "all tests rc=$?"; tail -8 gpurun_out/r14_tests.log
SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r14_bench.log 2>&1; tail -2 gpurun_out/r14_bench.log
