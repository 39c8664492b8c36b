#!/bin/bash
# evidence for profiles/: launch list of one fused bench step, full capture of the conv GEMM, stft launch list
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "stft" > gpurun_out/r15_stft_tests.log 2>&1
echo "stft tests rc=$?" ; tail -3 gpurun_out/r15_stft_tests.log
SE_BENCH_BATCH=64 timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/r15_bench_plain.log 2>&1 || exit 1
tail -c 600 gpurun_out/r15_bench_plain.log
SE_BENCH_BATCH=64 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/r15_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r15_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/r15_launches.csv
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/r15_stft_launches.csv python scripts/stft_bench.py > gpurun_out/r15_ncu_stft.log 2>&1
echo "ncu stft rc=$?"
SE_BENCH_BATCH=64 timeout 900 ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 24 -c 12 \
  -o gpurun_out/r15_gemm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/r15_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
