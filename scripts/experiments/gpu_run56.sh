#!/bin/bash
# final-state evidence: ncu launch list + full capture of the conv GEMM launches of one forward
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/r1i_bench_plain.log 2>&1 || { tail -5 gpurun_out/r1i_bench_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/r1i_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r1i_ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/r1i_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 32 -c 12 \
  -o gpurun_out/r1i_gemm_full -f python bench.py --steps 2 --warmup 3 > gpurun_out/r1i_ncu_full.log 2>&1
echo "ncu gemm rc=$?"; ls -la gpurun_out/r1i_gemm_full.ncu-rep
