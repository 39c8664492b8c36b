#!/bin/bash
# FFT engine: parity, microbench, ncu full capture of the two kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "stft" > gpurun_out/r27_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r27_tests.log
for b in 64 256 1024; do timeout 300 python scripts/stft_bench.py --batch $b 2>&1 | tail -1; done
timeout 300 python scripts/stft_bench.py --batch 256 --win 512 --hop 128 2>&1 | tail -1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fft_kernel -s 8 -c 4 \
  -o gpurun_out/r27_fft_full -f python scripts/stft_bench.py --batch 256 > gpurun_out/r27_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
