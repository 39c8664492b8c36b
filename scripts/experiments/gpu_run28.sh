#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "stft or dcunet" > gpurun_out/r32_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r32_tests.log
for b in 64 256 1024; do timeout 300 python scripts/stft_bench.py --batch $b 2>&1 | tail -1; done
