#!/bin/bash
# FFT engine for ConvSTFT / ConviSTFT: parity + microbench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "stft" > gpurun_out/r26_tests.log 2>&1
echo "tests rc=$?"; tail -25 gpurun_out/r26_tests.log
for b in 64 256 1024; do timeout 300 python scripts/stft_bench.py --batch $b 2>&1 | tail -3; done
timeout 300 python scripts/stft_bench.py --batch 256 --win 512 --hop 128 2>&1 | tail -3
