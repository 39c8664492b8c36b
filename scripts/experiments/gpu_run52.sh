#!/bin/bash
mkdir -p gpurun_out
for cfg in "180 0" "50 0" "50 1" "90 1" "180 1"; do
  set -- $cfg
  echo "== EPI_COL=$1 MMA_MODEL=$2"
  for m in dccrn dcunet20 frcrn; do
    SE_GEMM_EPI_COL=$1 SE_GEMM_MMA_MODEL=$2 timeout 300 python scripts/chain_layers.py --model $m > gpurun_out/r55_${m}_$1_$2.txt 2>&1
    grep TOTAL gpurun_out/r55_${m}_$1_$2.txt
  done
  grep -E "^ 8 |^10 " gpurun_out/r55_dccrn_$1_$2.txt
done
