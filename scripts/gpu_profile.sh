#!/bin/bash
# evidence for profiles/: launch list of the default bench, full ncu captures of the conv GEMM, LSTM, STFT / iSTFT FFT
# and mask kernels, per-layer chained conv timings, config-5 sweep.  usage: gpu_profile.sh <tag>
tag=${1:-r1e}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_bench_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_bench_plain.log; exit 1; }
tail -c 300 gpurun_out/${tag}_bench_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/${tag}_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 32 -c 16 \
  -o gpurun_out/${tag}_gemm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lstm_rec -s 4 -c 2 \
  -o gpurun_out/${tag}_lstm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_lstm.log 2>&1
echo "ncu lstm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fft_kernel|mask_kernel|cconv_first|col2im|cl2_from_lstm_rows' -s 18 -c 6 \
  -o gpurun_out/${tag}_hbm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_hbm.log 2>&1
echo "ncu hbm rc=$?"; ls -la gpurun_out/${tag}_*.ncu-rep
for m in dccrn dcunet20 frcrn; do timeout 300 python scripts/chain_layers.py --model $m; done > gpurun_out/${tag}_chain_layers.txt 2>&1
timeout 900 python scripts/stft_sweep.py --engine auto > gpurun_out/${tag}_stft_roundtrip_sweep.txt 2>&1
tail -2 gpurun_out/${tag}_stft_roundtrip_sweep.txt
