#!/bin/bash
# evidence for profiles/ (run under gpurun): launch list of the default bench; labelled full ncu capture of every conv
# launch of one DCCRN step at batch 64; full captures of the LSTM, STFT / iSTFT and mask kernels; per-layer chained conv
# timings; STFT sweep.  The .ncu-rep files are exported to raw CSV on the box and deleted (gpurun brings back <= 64 MiB);
# post-process here with scripts/ncu_layers.py / scripts/ncu_summary.py.     usage: scripts/gpu_profile.sh <tag>
tag=${1:-r2}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-others > gpurun_out/${tag}_bench_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_bench_plain.log; exit 1; }
tail -c 300 gpurun_out/${tag}_bench_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-others > gpurun_out/${tag}_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/${tag}_launches.csv
SE_GEMM_DEBUG=4 timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'cconv_gemm_kernel|cconv_first_kernel|cconv_small_cin|col2im_small' -o /tmp/${tag}_gemm_full -f \
  python scripts/ncu_conv_capture.py --model dccrn --batch 64 2> gpurun_out/${tag}_gemm_capture.log > /dev/null
echo "ncu gemm rc=$?"
ncu -i /tmp/${tag}_gemm_full.ncu-rep --page raw --csv > gpurun_out/${tag}_gemm_full_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'lstm_rec|fft_kernel|mask_kernel|cl2_from_lstm_rows' -o /tmp/${tag}_other_full -f \
  python scripts/ncu_conv_capture.py --model dccrn --batch 64 > gpurun_out/${tag}_ncu_other.log 2>&1
echo "ncu other rc=$?"
ncu -i /tmp/${tag}_other_full.ncu-rep --page raw --csv > gpurun_out/${tag}_other_full_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --profile-from-start off -o /tmp/${tag}_cublas_ref -f \
  python scripts/ncu_cublas_ref.py > gpurun_out/${tag}_ncu_cublas.log 2>&1
echo "ncu cublas rc=$?"
ncu -i /tmp/${tag}_cublas_ref.ncu-rep --page raw --csv > gpurun_out/${tag}_cublas_ref_raw.csv 2>/dev/null
ls -la /tmp/${tag}_*.ncu-rep gpurun_out/
timeout 120 python scripts/peak_tf32.py > gpurun_out/${tag}_peak_tf32.json 2>&1
timeout 300 python scripts/dropin_paths.py --model dccrn > gpurun_out/${tag}_dropin_paths.txt 2>&1
for m in dccrn dcunet20 frcrn; do timeout 300 python scripts/chain_layers.py --model $m; done > gpurun_out/${tag}_chain_layers.txt 2>&1
timeout 900 python scripts/stft_sweep.py --engine auto > gpurun_out/${tag}_stft_roundtrip_sweep.txt 2>&1
tail -2 gpurun_out/${tag}_stft_roundtrip_sweep.txt
