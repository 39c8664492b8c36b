#!/bin/bash
# evidence for profiles/: launch list of the fused bench, full captures of the conv GEMM and LSTM kernels
tag=${1:-r1c}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_bench_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_bench_plain.log; exit 1; }
tail -c 300 gpurun_out/${tag}_bench_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/${tag}_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 30 -c 15 \
  -o gpurun_out/${tag}_gemm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lstm_rec_kernel -s 4 -c 2 \
  -o gpurun_out/${tag}_lstm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_lstm.log 2>&1
echo "ncu lstm rc=$?"; ls -la gpurun_out/${tag}_*.ncu-rep
