#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stft_fft_kernel -s 8 -c 1 \
  -o gpurun_out/r41_fft_fwd -f python scripts/stft_bench.py --batch 256 > gpurun_out/r41_ncu.log 2>&1
echo "ncu rc=$?"
