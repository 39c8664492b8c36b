#!/bin/bash
# FFT engine v3: parity, microbench, ncu full capture of the two kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "stft" > gpurun_out/r29_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r29_tests.log
for b in 64 256 1024; do timeout 300 python scripts/stft_bench.py --batch $b 2>&1 | tail -1; done
timeout 300 python scripts/stft_bench.py --batch 256 --win 512 --hop 128 2>&1 | tail -1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stft_fft_kernel -s 8 -c 1 \
  -o gpurun_out/r29_fft_fwd -f python scripts/stft_bench.py --batch 256 > gpurun_out/r29_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:istft_fft_kernel -s 8 -c 1 \
  -o gpurun_out/r29_fft_inv -f python scripts/stft_bench.py --batch 256 > gpurun_out/r29_ncu2.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/r29*.ncu-rep
